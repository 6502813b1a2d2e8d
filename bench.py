#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native SL hot path (one JSON line on stdout).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload W]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (BASELINE.json configs):
  encode  configs[1]: 16384 synthetic random-playout 19x19 positions per step through the feature
          encoder; metric = positions/s; roofline = HBM (18,054 algorithmic bytes per position).
  train   configs[2]/[3]: SL policy net (13 convs, 192 filters, 48 planes) bf16 training step,
          batch 256 per GPU; metric = positions/s; roofline = tensor (8.2380 GFLOP per position).
`value` is device-timed (CUDA events on the launching stream, barrier + synchronize on both
sides, max over ranks) with inputs resident in HBM; `e2e` goes through the public Python API with
pinned HOST buffers and counts the copies.  `--impl reference` times the CPU oracle port of the
reference's algorithm on the host cores (the reference's own Keras/Theano stack is not installable
here; its Python encoder cannot travel to the GPU box).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FEATS46 = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
           "liberties_after", "sensibleness", "zeros"]
ENC_BYTES_PER_POSITION = 48 * 361 + 361 + 361 + 4  # planes out + board + age + player/ko
ENC_WARP_INST_PER_POSITION_48 = (365413068 + 947008922) / 16384.0   # profiles/r2_encode_ncu_summary.json


def encode_config(n, feats, planes):
    return {"workload": "feature-plane encoder, %d synthetic random-playout 19x19 positions/step, %d planes (%s)"
            % (n, planes, ",".join(feats)),
            "positions_per_step": n, "planes": planes,
            "l2": "4 rotating input sets; each step streams %d MB of output planes through the 126 MB L2"
            % (n * planes * 361 // 2 ** 20)}


def train_config(batch, dataset=4096):
    return {"workload": "AlphaGo SL policy net (13 convs: 5x5 + 11x 3x3 + 1x1, 192 filters, 48 planes) bf16 training step, "
                        "batch %d per GPU, device-side D4 augmentation, SGD lr decay" % batch,
            "per_gpu_batch": batch, "planes": 48,
            "l2": "each step touches ~1 GB of activations/gradients (> 126 MB L2); dataset of %d positions rotates" % dataset,
            "data": "planes from the CUDA encoder on random-playout positions, labels uniform over legal moves"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tensor_burst": d["bf16_tflops"], "tensor_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm": 6650.0, "tensor_burst": 1590.0, "tensor_sustained": 1400.0, "source": "fallback"}


class ClockSampler(object):
    """nvidia-smi sampler running during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.idx = device_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = vis.split(",")[self.idx] if vis else str(self.idx)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", phys, "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if sm:
            # samples "under load": power above 40% of the maximum seen, else all of them
            lim = 0.4 * max(pw) if pw else 0
            load = sorted(s for s, p in zip(sm, pw) if p >= lim) or sorted(sm)
            out.update(sm_mhz=load[len(load) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(pw))
        return out


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ------------------------------------------------------------------------------------------ encode
class EncodeWorkload(object):
    name = "encode"
    metric = "encode_positions_per_sec"
    unit = "positions/s"
    dtype = "u8"

    def __init__(self, n_positions, feats, rank):
        import numpy as np
        import torch
        from alphago_b200 import go
        from alphago_b200.preprocessing import Preprocess
        self.torch, self.np = torch, np
        self.n = n_positions
        self.feats = feats
        self.pp = Preprocess(feats)
        self.size = 19
        b, a, pl, ko = go.playout_positions(1000 + rank, n_positions, size=19, max_moves=400, per_game=16)
        # four input variants (board symmetries are again legal positions) so consecutive steps never
        # re-read the same input lines; each step also writes ~n*17 KB of planes through L2
        self.host = []
        for v in range(4):
            bb = b.reshape(-1, 19, 19); aa = a.reshape(-1, 19, 19)
            kx, ky = np.divmod(np.maximum(ko, 0), 19)
            if v & 1:
                bb, aa, kx, ky = bb.transpose(0, 2, 1), aa.transpose(0, 2, 1), ky, kx
            if v & 2:
                bb, aa, kx = bb[:, ::-1, :], aa[:, ::-1, :], 18 - kx
            kk = np.where(ko >= 0, kx * 19 + ky, -1).astype(np.int16)
            self.host.append((np.ascontiguousarray(bb).reshape(-1, 361), np.ascontiguousarray(aa).reshape(-1, 361), pl, kk))
        self.dev = [tuple(torch.from_numpy(x).cuda() for x in h) for h in self.host]
        self.out = torch.empty((n_positions, self.pp.output_dim, 19, 19), dtype=torch.uint8, device="cuda")
        self.i = 0
        self.units_per_step = n_positions
        self.launches_per_step = 1
        self.pinned = None

    def step(self):
        d = self.dev[self.i & 3]
        self.i += 1
        self.pp.encode_packed(d[0], d[1], d[2], d[3], self.size, out=self.out)

    def e2e_setup(self):
        t = self.torch
        self.pinned = [tuple(t.from_numpy(x).pin_memory() for x in h) for h in self.host]
        self.pin_out = t.empty(self.out.shape, dtype=t.uint8).pin_memory()
        self.h2d = sum(x.numel() * x.element_size() for x in self.pinned[0])
        self.d2h = self.pin_out.numel()

    def e2e_step(self):
        h = self.pinned[self.i & 3]
        self.i += 1
        d = [x.cuda(non_blocking=True) for x in h]
        self.pp.encode_packed(d[0], d[1], d[2], d[3], self.size, out=self.out)
        self.pin_out.copy_(self.out, non_blocking=True)
        self.torch.cuda.current_stream().synchronize()
        return int(self.pin_out[0, 3, 0, 0])

    def roofline(self, ms_per_step, pk):
        ach = self.n * ENC_BYTES_PER_POSITION / (ms_per_step * 1e-3) / 1e9
        out = {"bound": "hbm", "achieved": round(ach, 2), "peak": pk["hbm"], "unit": "GB/s",
               "frac": round(ach / pk["hbm"], 5), "traffic": None, "kernel": "ag::enc::encode_kernel + ag::enc::ladder_resolve_kernel",
               "peak_source": pk["source"] + " (copy bandwidth, kernel timed inside a long step)",
               "algorithmic_bytes_per_launch": self.n * ENC_BYTES_PER_POSITION}
        tp = os.path.join(ROOT, "profiles", "r2_traffic.json")
        if os.path.exists(tp) and self.n == 16384 and self.pp.output_dim == 48:
            t = json.load(open(tp))
            out["traffic"] = t["encode_kernel"]["dram_bytes_per_launch"] + t["ladder_resolve_kernel"]["dram_bytes_per_launch"]
        # The encoder is integer work: what bounds it in practice is instruction issue, not bandwidth.  Warp instructions per
        # position from the committed ncu capture (profiles/r2_encode_ncu_summary.json: smsp__inst_executed.sum / 16384
        # positions, 48 planes) against 148 SMs x 4 schedulers x 1 instruction per cycle at the SM clock seen under this load.
        if self.pp.output_dim == 48:
            inst = ENC_WARP_INST_PER_POSITION_48
            clk = 1.9e9
            floor_ms = self.n * inst / (148 * 4 * clk) * 1e3
            out["issue_roofline"] = {"warp_instructions_per_position": inst, "sm_clock_hz": clk, "floor_ms": round(floor_ms, 4),
                                     "frac": round(floor_ms / ms_per_step, 4),
                                     "note": "encode_kernel 22.3 k + ladder_resolve_kernel 57.8 k warp instructions per position"}
        return out

    def config(self):
        return encode_config(self.n, self.feats, self.pp.output_dim)

    def cpu_baseline(self, sample=2048):
        from oracle import go_oracle as og
        cores = os.cpu_count() or 1
        h = self.host[0]
        s = min(sample, self.n)
        og.encode_batch(h[0][:64], h[1][:64], h[2][:64], h[3][:64], 19, self.feats, threads=cores)
        t = time.time()
        reps = 0
        while time.time() - t < 5.0 or reps < 1:
            og.encode_batch(h[0][:s], h[1][:s], h[2][:s], h[3][:s], 19, self.feats, threads=cores)
            reps += 1
        dt = time.time() - t
        return {"value": round(reps * s / dt, 1), "unit": self.unit, "cores": cores, "kind": "port",
                "sample": "%d x %d of the step's positions, C oracle (oracle/go_oracle.c) with OpenMP over all host cores, %.2f s"
                          % (reps, s, dt)}


# ------------------------------------------------------------------------------------------ train
TRAIN_FLOP_PER_POSITION = 8.2380e9   # fwd 2.8015 + wgrad 2.8015 + dgrad 2.6352 GFLOP (192 filters, 48 planes)
FWD3x3_FLOP_PER_POSITION = 2 * 361 * 9 * 192 * 192  # one 3x3 192->192 layer, algorithmic (no padding flops)
FWD_FLOP_PER_POSITION = 2 * 361 * (25 * 48 * 192 + 11 * 9 * 192 * 192)    # the 12 KxK convs forward (2.8014 GFLOP)
DGRAD_FLOP_PER_POSITION = 2 * 361 * 11 * 9 * 192 * 192                   # data gradients of layers 11..1 (2.6351 GFLOP)
TRAIN_FEATS = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
               "ladder_capture", "ladder_escape", "sensibleness", "zeros"]  # the default 48 planes (preprocessing.py:211-214)


class TrainWorkload(object):
    name = "train"
    metric = "sl_policy_train_positions_per_sec"
    unit = "positions/s"
    dtype = "bf16"

    def __init__(self, batch, rank, dataset=4096):
        import numpy as np
        import torch
        from alphago_b200 import go, nn
        from alphago_b200.preprocessing import Preprocess
        from alphago_b200.training import SLTrainer
        self.torch, self.np = torch, np
        self.batch = batch
        rng = np.random.RandomState(77 + rank)
        # synthetic dataset: random-playout positions encoded by the CUDA encoder, labels uniform over legal moves
        b, a, pl, ko = go.playout_positions(500 + rank, dataset, size=19, max_moves=400, per_game=16)
        pp = Preprocess(TRAIN_FEATS)
        self.states = pp.encode_packed(*[torch.from_numpy(x).cuda() for x in (b, a, pl, ko)], 19)
        legal = self.states[:, 20:28].sum(1).reshape(dataset, 361).cpu().numpy() > 0
        acts = np.zeros((dataset, 2), np.uint8)
        for i in range(dataset):
            mv = np.flatnonzero(legal[i])
            m = int(mv[rng.randint(len(mv))]) if len(mv) else 0
            acts[i] = divmod(m, 19)
        self.h_states = self.states.cpu().numpy()
        self.h_actions = acts
        self.actions = torch.from_numpy(acts).cuda()
        self.net = nn.ConvNet(19, 48, [192] * 12, [5] + [3] * 11, max_batch=batch)
        self.net.init_uniform(seed=1)  # identical weights on every rank
        self.trainer = SLTrainer(self.net, lr=0.003, decay=0.0001, momentum=0.0)
        self.dataset = dataset
        self.rng = rng
        self.i = 0
        self.units_per_step = batch
        self.launches_per_step = None

    def _next(self):
        idx = (self.np.arange(self.batch) + self.i * self.batch) % self.dataset
        self.i += 1
        return idx, self.rng.randint(0, 8, self.batch)

    def step(self):
        idx, tf = self._next()
        x, y = self.trainer.assemble(self.states, self.actions, idx, tf)
        self.trainer.step(x, y)

    def e2e_setup(self):
        from alphago_b200.training import HostFeeder
        self.feeder = HostFeeder(self.trainer, self.h_states, self.h_actions, self.batch)
        self.h2d = self.feeder.h2d_bytes + self.batch * 9   # + the gather index (int64) and symmetry (uint8) vectors
        self.d2h = self.feeder.d2h_bytes

    def e2e_step(self):
        # the public training path with a HOST dataset (training.HostFeeder): gather into pinned memory, copy, augment +
        # step on the device, read the running loss back; the copy of batch k overlaps the kernels of step k-1
        idx, tf = self._next()
        return self.feeder.step(idx, tf)

    def e2e_finish(self):
        return self.feeder.flush()

    def kernel_rooflines(self, pk):
        """per-launch device time of the step's GEMM kernels, CUDA events on the launching stream: the chained forward
        stack and data-gradient chain (one launch each, as the step runs them) and the kernels of a middle 3x3 layer"""
        t = self.torch
        out = {}
        f3 = self.batch * FWD3x3_FLOP_PER_POSITION
        cases = [(3, "conv_chain_kernel(forward stack, 12 layers)", self.batch * FWD_FLOP_PER_POSITION),
                 (4, "conv_chain_kernel(data gradients, 11 layers)", self.batch * DGRAD_FLOP_PER_POSITION),
                 (0, "conv_gemm_win_kernel(fwd 3x3, one layer)", f3), (1, "conv_gemm_win_kernel(dgrad 3x3, one layer)", f3),
                 (5, "wgrad_chain_kernel(weight gradients of the 11 3x3 layers, split reduction included)", 11 * f3),
                 (2, "wgrad_kernel+reduce(3x3, one layer, stand-alone)", f3)]
        for mode, name, flop in cases:
            # "timed alone" means burst conditions, which is what the burst peak it is divided by was measured under: let the
            # power governor settle after the long step loop (under sustained load the SM clock sits near 1.65 GHz)
            t.cuda.synchronize()
            time.sleep(1.0)
            try:
                for _ in range(3):
                    self.net.run_layer(6, mode, self.batch)
            except Exception:
                continue   # the net runs unchained (AG_CONV_CHAIN=0)
            e0, e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
            reps = 20
            e0.record()
            for _ in range(reps):
                self.net.run_layer(6, mode, self.batch)
            e1.record()
            t.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            out[name] = {"ms": round(ms, 4), "tflops": round(flop / (ms * 1e-3) / 1e12, 1),
                         "frac_of_burst_peak": round(flop / (ms * 1e-3) / 1e12 / pk["tensor_burst"], 4), "flop": flop}
        return out

    def in_step_durations(self, steps=10):
        """per-kernel durations INSIDE the running step (CUPTI through torch.profiler, after the timed region): the
        same kernels timed alone run at burst clocks with a warm L2, inside the step they follow each other"""
        t = self.torch
        try:
            from torch.profiler import ProfilerActivity, profile
            for _ in range(3):
                self.step()
            t.cuda.synchronize()
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                for _ in range(steps):
                    self.step()
                t.cuda.synchronize()
            acc = {}
            evs = [e for e in prof.events() if e.time_range is not None and "cuda" in str(e.device_type).lower()]
            for e in sorted(evs, key=lambda e: e.time_range.start):
                name = e.name.replace("(anonymous namespace)::", "").split("(")[0].split("::")[-1].split("<")[0].strip()
                if name.startswith("Memcpy") or name.startswith("Memset"):
                    continue
                acc.setdefault(name, []).append(e.time_range.end - e.time_range.start)
            out = {}
            for k, v in acc.items():
                lps = len(v) // steps if len(v) % steps == 0 else 0
                out[k] = {"us_per_step": round(sum(v) / steps, 2), "launches_per_step": round(len(v) / steps, 2)}
                if 1 <= lps <= 4:   # the i-th launch of this kernel in a step, averaged over the steps
                    out[k]["us_each"] = [round(sum(v[o::lps]) / steps, 2) for o in range(lps)]
            return out
        except Exception as e:  # the profiler is an aid: the bench line does not depend on it
            return {"unavailable": str(e)[:120]}

    def timeline(self, path, steps=4):
        """per-stream GPU timeline of one training step on this rank (CUPTI through torch.profiler's chrome trace): every kernel
        with its stream, start and duration, NCCL's kernels included, and the time between the end of the last backward kernel
        and the start of the SGD kernel -- what the gradient exchange leaves exposed"""
        import tempfile
        t = self.torch
        from torch.profiler import ProfilerActivity, profile
        for _ in range(3):
            self.step()
        t.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(steps):
                self.step()
            t.cuda.synchronize()
        fd, tmp = tempfile.mkstemp(suffix=".json")
        os.close(fd)
        prof.export_chrome_trace(tmp)
        ev = [e for e in json.load(open(tmp)).get("traceEvents", []) if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
        os.unlink(tmp)
        ev.sort(key=lambda e: e["ts"])
        sgd = [i for i, e in enumerate(ev) if "sgd_all" in e["name"]]
        if len(sgd) < 3:
            return None
        a, z = sgd[-3] + 1, sgd[-2] + 1          # one whole step: after an SGD kernel up to and including the next
        rows = ev[a:z]
        t0 = rows[0]["ts"]
        lines = ["# one training step on rank %s of %s: kernels in start order; times in us from the step's first kernel" % (
            os.environ.get("RANK", "0"), os.environ.get("WORLD_SIZE", "1")), "# %-8s %10s %10s  %s" % ("stream", "start", "dur", "kernel")]
        last_bwd_end, first_nccl, last_nccl_end, sgd_start = 0.0, None, 0.0, None
        for e in rows:
            name = e["name"].replace("(anonymous namespace)::", "").split("(")[0]
            st = e.get("args", {}).get("stream", "?")
            lines.append("  %-8s %10.1f %10.1f  %s" % (st, e["ts"] - t0, e.get("dur", 0.0), name[-70:]))
            end = e["ts"] - t0 + e.get("dur", 0.0)
            if "nccl" in name.lower():
                first_nccl = e["ts"] - t0 if first_nccl is None else first_nccl
                last_nccl_end = max(last_nccl_end, end)
            elif "sgd_all" in name:
                sgd_start = e["ts"] - t0
            elif "wgrad" in name or "conv_chain" in name:
                last_bwd_end = max(last_bwd_end, end)
        lines.append("# step span %.1f us; last backward kernel ends at %.1f" % (rows[-1]["ts"] - t0 + rows[-1].get("dur", 0.0), last_bwd_end))
        if first_nccl is not None and sgd_start is not None:
            lines.append("# first NCCL kernel starts at %.1f (%.1f us before the backward pass ends), last ends at %.1f; SGD starts at %.1f: "
                         "%.1f us between the end of backward and the update" % (first_nccl, last_bwd_end - first_nccl, last_nccl_end, sgd_start,
                                                                               sgd_start - last_bwd_end))
        open(path, "w").write("\n".join(lines) + "\n")
        return path

    def sustained(self, seconds=2.2, ms_per_step=1.7):
        """the same step for >= `seconds` back to back with its own clock samples: the power governor settles (the SM clock
        drops to ~1.65 GHz), which is what the sustained cuBLAS peak was measured under"""
        t = self.torch
        steps = int(seconds / (ms_per_step * 1e-3)) + 1
        sampler = ClockSampler(t.cuda.current_device())
        sampler.start()
        e0, e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
        t.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            self.step()
        e1.record()
        t.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        return steps, ms, sampler.stop()

    def roofline(self, ms_per_step, pk, steps=100, world=1, sustained=None):
        ach = self.batch * TRAIN_FLOP_PER_POSITION / (ms_per_step * 1e-3) / 1e12
        k = self.kernel_rooflines(pk)
        name = next(iter(k))   # the forward chain when the net runs chained, else the single-layer forward kernel
        dom = k[name]
        traffic = None
        for tp in ("r2_traffic.json", "r1_traffic.json"):
            tp = os.path.join(ROOT, "profiles", tp)
            if os.path.exists(tp):  # DRAM bytes per launch of this kernel from the committed ncu --set full capture
                traffic = json.load(open(tp)).get(name.split("(")[0], {}).get("dram_bytes_per_launch")
                break
        # the whole step against the peak that matches its timed region: a region shorter than a second runs at burst clocks
        region_s = steps * ms_per_step * 1e-3
        kind = "tensor_burst" if region_s < 1.0 else "tensor_sustained"
        out = {"bound": "tensor", "achieved": dom["tflops"], "peak": pk["tensor_burst"], "unit": "TFLOP/s",
               "frac": dom["frac_of_burst_peak"], "traffic": traffic,
               "kernel": "ag::" + name + ", timed alone (20 launches back to back)",
               "peak_source": pk["source"] + " (cuBLAS bf16 burst, kernel timed alone)",
               "algorithmic_flops_per_launch": dom["flop"],
               "whole_step": {"achieved": round(ach, 1), "peak": pk[kind], "frac": round(ach / pk[kind], 4), "unit": "TFLOP/s",
                              "peak_kind": kind, "timed_region_s": round(region_s, 3),
                              "note": "8.2380 GFLOP/position over the whole step (all %d launches) vs the peak measured under "
                                      "the same conditions as the timed region" % (self.launches_per_step or 0)},
               "kernels": k}
        if sustained is not None:
            n, ms, clocks = sustained
            a = self.batch * TRAIN_FLOP_PER_POSITION * n / (ms * 1e-3) / 1e12
            out["whole_step_sustained"] = {"achieved": round(a, 1), "peak": pk["tensor_sustained"], "unit": "TFLOP/s",
                                           "frac": round(a / pk["tensor_sustained"], 4), "steps": n, "seconds": round(ms * 1e-3, 3),
                                           "positions_per_sec_per_gpu": round(self.batch * n / (ms * 1e-3), 1), "clocks": clocks}
        return out

    def config(self):
        return train_config(self.batch, self.dataset)

    def cpu_baseline(self):
        import torch
        from oracle import nn_oracle as no
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        params = no.make_params(48, [192] * 12, [5] + [3] * 11, seed=1)
        bs = self.batch   # whole minibatches of the step's own data
        x = self.h_states[:bs]
        y = self.h_actions[:bs, 0].astype("int64") * 19 + self.h_actions[:bs, 1]
        no.loss_and_grads(params, x[:8], y[:8])
        t = time.time()
        steps = 0
        while time.time() - t < 10 or steps < 1:
            _, _, grads, _ = no.loss_and_grads(params, x, y)
            params, _ = no.sgd_update(params, grads, 0.003)
            steps += 1
        dt = time.time() - t
        return {"value": round(steps * bs / dt, 2), "unit": self.unit, "cores": cores, "kind": "port",
                "sample": "%d steps at batch %d of the same architecture, fp32 torch-CPU restatement (oracle/nn_oracle.py), %.1f s; "
                          "the reference's Keras 1.0/Theano 0.8 path is unavailable on this box" % (steps, bs, dt)}


def reference_arm(args, rank, workload):
    """The reference arm: the reference's own CPU algorithm for the path on the host cores, same metric/config as our
    arm.  The reference's Python (8 ms/position encoder) cannot travel to the GPU box and its Keras 1.0/Theano 0.8
    stack is not installable, so this is the oracle port: oracle/go_oracle.c (OpenMP over all cores) for `encode`,
    the fp32 torch-CPU restatement of the network (oracle/nn_oracle.py, all cores) for `train`.  Each step is the
    whole step of the repo arm (256 positions of training, 16384 of encoding: about 1-2 s of host time each)."""
    if rank != 0:
        return
    import numpy as np
    from oracle import go_oracle as og
    cores = os.cpu_count() or 1
    if workload == "encode":
        feats = FEATS46 if args.planes == 46 else og.DEFAULT_FEATURES
        sample = args.positions   # the whole step: same configuration as the repo arm
        b, a, pl, ko = og.playout_positions(1000, sample, size=19, max_moves=400, per_game=16)
        for _ in range(args.warmup):
            og.encode_batch(b[:256], a[:256], pl[:256], ko[:256], 19, feats, threads=cores)
        t = time.time()
        for _ in range(args.steps):
            og.encode_batch(b, a, pl, ko, 19, feats, threads=cores)
        dt = time.time() - t
        metric, dtype = EncodeWorkload.metric, "u8"
        cfg = encode_config(sample, feats, og.output_dim(feats))
        how = "all %d positions of the step, C oracle (oracle/go_oracle.c) with OpenMP over all host cores" % sample
    else:
        import torch
        from oracle import nn_oracle as no
        torch.set_num_threads(cores)
        sample = args.batch       # the whole minibatch: same configuration as the repo arm
        params = no.make_params(48, [192] * 12, [5] + [3] * 11, seed=1)
        rng = np.random.RandomState(0)
        b, a, pl, ko = og.playout_positions(500, sample, size=19, max_moves=400, per_game=16)
        x = og.encode_batch(b, a, pl, ko, 19, og.DEFAULT_FEATURES, threads=cores)
        y = rng.randint(0, 361, sample)
        for _ in range(min(args.warmup, 2)):
            no.loss_and_grads(params, x[:8], y[:8])
        t = time.time()
        for _ in range(args.steps):
            _, _, grads, _ = no.loss_and_grads(params, x, y)
            params, _ = no.sgd_update(params, grads, 0.003)
        dt = time.time() - t
        metric, dtype = TrainWorkload.metric, "f32"
        cfg = train_config(sample)
        how = ("all %d positions of the step, fp32 torch-CPU restatement of the same net and SGD step (oracle/nn_oracle.py), "
               "all host cores; the reference's Keras 1.0 / Theano 0.8 stack is not installable here" % sample)
    v = round(args.steps * sample / dt, 2)
    line = {"impl": "reference", "metric": metric, "value": v, "unit": "positions/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": v, "unit": "positions/s", "cores": cores, "kind": "port", "sample": how},
            "e2e": {"value": v, "unit": "positions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["encode", "train"])
    ap.add_argument("--positions", type=int, default=16384)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--planes", type=int, default=48, choices=[46, 48])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-encoder", action="store_true")
    ap.add_argument("--no-sustained", action="store_true", help="skip the 2 s sustained run and the in-step kernel timeline")
    ap.add_argument("--timeline", default=None, help="train workload: write rank 0's per-stream kernel timeline of one step to this file")
    args = ap.parse_args()
    rank, world, local = dist_env()
    workload = args.workload or "train"
    if args.steps is None:
        args.steps = (200 if workload == "encode" else 100) if args.impl == "ours" else 5
    if args.warmup is None:
        args.warmup = 5 if args.impl == "ours" else 1
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        reference_arm(args, rank, workload)
        return

    # libraries (NCCL's version banner, torchrun notices) write to stdout: keep fd 1 for the one JSON line
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        os.write(json_fd, (json.dumps(line) + "\n").encode())

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from alphago_b200 import _lib

    feats = FEATS46 if args.planes == 46 else TRAIN_FEATS
    if workload == "train":
        wl = TrainWorkload(args.batch, rank)
    else:
        wl = EncodeWorkload(args.positions, feats, rank)
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        wl.step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    l0 = _lib.lib().ag_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        wl.step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = _lib.lib().ag_launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * wl.units_per_step * args.steps / (ms * 1e-3)

    # end to end through the public API with host buffers
    wl.e2e_setup()
    ke = max(3, min(args.steps, 50))
    wl.e2e_step()
    barrier()
    t0 = time.time()
    for _ in range(ke):
        wl.e2e_step()
    if hasattr(wl, "e2e_finish"):
        wl.e2e_finish()   # the last step's result is read back inside the timed region too
    torch.cuda.synchronize()
    te = time.time() - t0
    if world > 1:
        t = torch.tensor([te], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        te = float(t.item())
    e2e = {"value": round(world * wl.units_per_step * ke / te, 1), "unit": wl.unit, "h2d_bytes_per_step": wl.h2d,
           "d2h_bytes_per_step": wl.d2h, "steps": ke}

    sustained = in_step = None
    if workload == "train" and not args.no_sustained:
        # every rank runs these (the step holds a collective when world > 1); only rank 0 reports them
        wl.launches_per_step = int(launches) // max(1, args.steps)
        sustained = wl.sustained(2.2, ms_per_step)
        barrier()
        in_step = wl.in_step_durations()
        barrier()
    if workload == "train" and args.timeline:
        try:   # every rank profiles (the step holds a collective); rank 0 writes
            wl.timeline(args.timeline if rank == 0 else os.devnull)
        except Exception as e:
            sys.stderr.write("timeline failed: %s\n" % e)
        barrier()

    if rank == 0:
        if workload == "train":
            roof = wl.roofline(ms_per_step, pk, steps=args.steps, world=world, sustained=sustained)
            if in_step is not None:
                roof["in_step"] = in_step
                for kname, kv in roof["kernels"].items():   # the in-step figure beside the timed-alone one
                    base = kname.split("(")[0].split("+")[0]
                    each = in_step.get(base, {}).get("us_each") if isinstance(in_step.get(base), dict) else None
                    if not each or "one layer" in kname:
                        continue
                    which = 1 if ("data gradients" in kname and len(each) > 1) else 0
                    kv["in_step_ms"] = round(each[which] / 1e3, 4)
                dom = next(iter(roof["kernels"].values()))
                if "in_step_ms" in dom:
                    a = dom["flop"] / (dom["in_step_ms"] * 1e-3) / 1e12
                    roof["in_step_frac"] = {"achieved": round(a, 1), "ms": dom["in_step_ms"],
                                            "frac_of_burst_peak": round(a / pk["tensor_burst"], 4),
                                            "frac_of_sustained_peak": round(a / pk["tensor_sustained"], 4),
                                            "note": "the dominant kernel's duration inside the running step (CUPTI); small kernels "
                                                    "launched with programmatic dependent launch start early and their in-step "
                                                    "durations include the wait for the kernel in front"}
        else:
            roof = wl.roofline(ms_per_step, pk)
        line = {"metric": wl.metric, "value": round(value, 1), "unit": wl.unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": wl.dtype, "data": "synthetic", "config": wl.config(),
                "roofline": roof, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = wl.cpu_baseline()
        if workload == "train" and not args.no_encoder:
            # the metric's second half: 48-plane encode positions/s on this GPU (configs[1]), same timing rules
            ew = EncodeWorkload(args.positions, feats, rank)
            for _ in range(3):
                ew.step()
            torch.cuda.synchronize()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record()
            for _ in range(50):
                ew.step()
            f1.record()
            torch.cuda.synchronize()
            ems = f0.elapsed_time(f1) / 50
            line["encoder"] = {"metric": ew.metric, "value": round(ew.n / (ems * 1e-3), 1), "unit": ew.unit, "ms_per_step": round(ems, 4),
                               "config": ew.config(), "roofline": ew.roofline(ems, pk)}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
