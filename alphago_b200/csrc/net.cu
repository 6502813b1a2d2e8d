// The conv-stack network object behind CNNPolicy / CNNValue (AlphaGo/models/policy.py:93-156,
// AlphaGo/models/value.py:12-28) and its SL training step (supervised_policy_trainer.py:199-213):
// device buffers, weight import/export, forward, fused softmax/cross-entropy head, backward and the
// SGD update.  The GEMM-shaped work runs in conv.cu / wgrad.cu (tcgen05); this file holds the
// memory-bound kernels around them and the orchestration.  Everything is allocated in
// ag_net_create; no entry point allocates or synchronises afterwards (except the host<->device
// weight/debug copies, which say so).
#include <math.h>
#include <string.h>

#include <stdlib.h>
#include <vector>

#include "common.cuh"
#include "conv.cuh"
#include "wgrad.cuh"

namespace ag {

typedef __nv_bfloat16 bf16;

static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// ------------------------------------------------------------------------------------ kernels

// uint8 planes [n][C][S][S] (the encoder's / HDF5 layout) -> bf16 padded-pixel rows [n*img_stride][64k]
// for the first conv layer; channels >= C and pad slots stay zero (buffer zero-filled once).
__global__ void hold_sm_kernel(long long ns) {
    extern __shared__ uint8_t hold_smem[];
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do {
        __nanosleep(1000);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    } while ((long long)(t - t0) < ns);
    if (ns < 0) hold_smem[threadIdx.x] = 0;
}

__global__ void pack_input_kernel(const uint8_t* __restrict__ planes, int n, int C, int S, PixelLayout L, int ld,
                                  bf16* __restrict__ out) {
    pdl_trigger();
    pdl_wait();
    // one thread per (image, point, 8-channel group): reads 8 planes' bytes, writes 16 B
    const int groups = ld / 8;
    const int pts = S * S;
    size_t total = (size_t)n * pts * groups;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int p = (int)(i % pts);
        size_t r = i / pts;
        int g = (int)(r % groups);
        int img = (int)(r / groups);
        int x = p / S, y = p - x * S;
        uint4 o;
        bf16* ob = reinterpret_cast<bf16*>(&o);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            int c = g * 8 + j;
            float v = c < C ? (float)planes[((size_t)img * C + c) * pts + p] : 0.f;
            ob[j] = __float2bfloat16(v);
        }
        size_t row = (size_t)img * L.img_stride + (size_t)x * L.row_stride + y;
        *reinterpret_cast<uint4*>(out + row * ld + g * 8) = o;
    }
}

// Policy head, forward (+ backward when labels are given).  One CTA per image.
//   logits[p] = sum_c X[p][c] * w[c] + b (+ pos_bias[p])          policy.py:144-150 (1x1 conv, Flatten)
//   probs = softmax(logits)                                        policy.py:154
//   loss  = -log(clip(probs[label], 1e-7, 1-1e-7)), acc = argmax == label   spt.py:200 (Keras CE / accuracy)
//   dlogit = (probs - onehot) * inv_batch; G[p][c] = dlogit[p] * w[c] * (X[p][c] > 0)
//   per-image partials of dw[c], db, dpos[p] go to `partial` (reduced by head_reduce_kernel)
constexpr int kHeadThreads = 384;
__global__ void __launch_bounds__(kHeadThreads)
policy_head_kernel(const bf16* __restrict__ X, int ld, PixelLayout L, const float* __restrict__ w, const float* __restrict__ bias,
                   const float* __restrict__ pos_bias, float* __restrict__ probs, const int32_t* __restrict__ labels,
                   float inv_batch, bf16* __restrict__ G, float* __restrict__ partial, float* __restrict__ loss_acc,
                   float* __restrict__ dlogits_out, int loss_kind) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s_logit[368];
    __shared__ float s_w[256];
    __shared__ float s_red[kHeadThreads / 32];
    __shared__ float s_dw[kHeadThreads / 32][256];  // per-warp partials, summed in warp order (deterministic)
    const int img = blockIdx.x, S = L.size, pts = S * S;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = kHeadThreads / 32;
    for (int c = threadIdx.x; c < 256; c += kHeadThreads) s_w[c] = c < ld ? w[c] : 0.f;
    __syncthreads();
    float wl[8];   // this lane's eight channel weights: loop-invariant, kept out of the row loops' shared-memory traffic
#pragma unroll
    for (int j = 0; j < 8; j++) wl[j] = s_w[(lane * 8 + j) & 255];
    const int groups = ld / 8;  // 16-byte groups per row; lanes >= groups idle (ld <= 256)
    const float b0 = bias[0];
    constexpr int U = 5;  // rows in flight per warp (measured: 8 with two CTAs per SM forced was slower, 40.8 vs 38.2 us): the loop is bound by load latency, not bandwidth
    for (int p0 = warp; p0 < pts; p0 += nwarps * U) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int p = p0 + u * nwarps;
            v[u] = make_uint4(0, 0, 0, 0);
            if (p < pts && lane < groups) {
                const int x = p / S, y = p - x * S;
                v[u] = *reinterpret_cast<const uint4*>(X + ((size_t)img * L.img_stride + (size_t)x * L.row_stride + y) * ld + lane * 8);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int p = p0 + u * nwarps;
            const bf16* vb = reinterpret_cast<const bf16*>(&v[u]);
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 8; j++) acc += __bfloat162float(vb[j]) * wl[j];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0 && p < pts) s_logit[p] = acc + b0 + (pos_bias ? pos_bias[p] : 0.f);
        }
    }
    __syncthreads();
    // softmax over pts (block reductions)
    float mx = -INFINITY;
    for (int p = threadIdx.x; p < pts; p += kHeadThreads) mx = fmaxf(mx, s_logit[p]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
    mx = s_red[0];
    for (int i = 1; i < nwarps; i++) mx = fmaxf(mx, s_red[i]);
    __syncthreads();
    float sum = 0.f;
    for (int p = threadIdx.x; p < pts; p += kHeadThreads) { float e = expf(s_logit[p] - mx); s_logit[p] = e; sum += e; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) s_red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int i = 0; i < nwarps; i++) sum += s_red[i];
    const float inv = 1.f / sum;
    for (int p = threadIdx.x; p < pts; p += kHeadThreads) {
        float pr = s_logit[p] * inv;
        s_logit[p] = pr;
        if (probs) probs[(size_t)img * pts + p] = pr;
    }
    __syncthreads();
    if (!labels) return;

    const int label = labels[img];
    if (warp == 0) {
        // accuracy: first maximum, as numpy/Theano argmax (ties go to the lowest index)
        int best = 0x7fffffff;
        float bv = -INFINITY;
        for (int p = lane; p < pts; p += 32) if (s_logit[p] > bv) { bv = s_logit[p]; best = p; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int ob = __shfl_xor_sync(0xffffffffu, best, o);
            if (ov > bv || (ov == bv && ob < best)) { bv = ov; best = ob; }
        }
        if (lane == 0) {
            if (loss_kind == 0) {
                float pl = fminf(fmaxf(s_logit[label], 1e-7f), 1.f - 1e-7f);
                atomicAdd(&loss_acc[0], -logf(pl));
            }
            atomicAdd(&loss_acc[1], best == label ? 1.f : 0.f);
        }
    }
    __syncthreads();
    // Keras' binary_crossentropy on the softmax output (reinforcement_policy_trainer.py:109): per sample the MEAN over the pts
    // outputs of -(y log p + (1-y) log(1-p)), p clipped to [1e-7, 1-1e-7] (no gradient where the clip is active).
    // With a_j = dL/dp_j the logit gradient is p_k (a_k - sum_j p_j a_j).
    float bce_s = 0.f;
    if (loss_kind == 1) {
        float ls = 0.f, ps = 0.f;
        for (int p = threadIdx.x; p < pts; p += kHeadThreads) {
            const float pr = s_logit[p], pc = fminf(fmaxf(pr, 1e-7f), 1.f - 1e-7f);
            const bool on = p == label, inside = pr > 1e-7f && pr < 1.f - 1e-7f;
            ls += on ? -logf(pc) : -logf(1.f - pc);
            const float a = inside ? (on ? -1.f / pc : 1.f / (1.f - pc)) : 0.f;
            ps += pr * a;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { ls += __shfl_xor_sync(0xffffffffu, ls, o); ps += __shfl_xor_sync(0xffffffffu, ps, o); }
        __shared__ float s_bce[2][kHeadThreads / 32];
        if (lane == 0) { s_bce[0][warp] = ls; s_bce[1][warp] = ps; }
        __syncthreads();
        ls = 0.f;
        for (int i = 0; i < nwarps; i++) { ls += s_bce[0][i]; bce_s += s_bce[1][i]; }
        if (threadIdx.x == 0) atomicAdd(&loss_acc[0], ls / pts);
    }
    // dlogit in place
    float dbl = 0.f;
    for (int p = threadIdx.x; p < pts; p += kHeadThreads) {
        float d;
        if (loss_kind == 0) d = (s_logit[p] - (p == label ? 1.f : 0.f)) * inv_batch;
        else {
            const float pr = s_logit[p], pc = fminf(fmaxf(pr, 1e-7f), 1.f - 1e-7f);
            const bool inside = pr > 1e-7f && pr < 1.f - 1e-7f;
            const float a = inside ? (p == label ? -1.f / pc : 1.f / (1.f - pc)) : 0.f;
            d = pr * (a - bce_s) * (inv_batch / pts);
        }
        s_logit[p] = d;
        dbl += d;
        if (dlogits_out) dlogits_out[(size_t)img * pts + p] = d;
        if (partial) partial[(size_t)img * (256 + 1 + 368) + 257 + p] = d;  // d(pos_bias)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dbl += __shfl_xor_sync(0xffffffffu, dbl, o);
    __syncthreads();
    if (lane == 0) s_red[warp] = dbl;
    __syncthreads();
    // G rows and dw partial
    float dw[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int p0 = warp; p0 < pts; p0 += nwarps * U) {
        uint4 v[U];
        size_t roff[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int p = p0 + u * nwarps;
            v[u] = make_uint4(0, 0, 0, 0);
            roff[u] = 0;
            if (p < pts && lane < groups) {
                const int x = p / S, y = p - x * S;
                roff[u] = ((size_t)img * L.img_stride + (size_t)x * L.row_stride + y) * ld + lane * 8;
                v[u] = *reinterpret_cast<const uint4*>(X + roff[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int p = p0 + u * nwarps;
            if (p < pts && lane < groups) {
                const float d = s_logit[p];
                const bf16* vb = reinterpret_cast<const bf16*>(&v[u]);
                uint4 o;
                bf16* ob = reinterpret_cast<bf16*>(&o);
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    float xv = __bfloat162float(vb[j]);
                    dw[j] += d * xv;
                    ob[j] = __float2bfloat16(xv > 0.f ? d * wl[j] : 0.f);
                }
                *reinterpret_cast<uint4*>(G + roff[u]) = o;
            }
        }
    }
    if (lane < groups) {
#pragma unroll
        for (int j = 0; j < 8; j++) s_dw[warp][lane * 8 + j] = dw[j];
    }
    __syncthreads();
    if (partial) {
        float* pp = partial + (size_t)img * (256 + 1 + 368);
        for (int c = threadIdx.x; c < ld; c += kHeadThreads) {
            float t = 0.f;
            for (int k = 0; k < nwarps; k++) t += s_dw[k][c];
            pp[c] = t;
        }
        if (threadIdx.x == 0) {
            float t = 0.f;
            for (int i = 0; i < nwarps; i++) t += s_red[i];
            pp[256] = t;
        }
    }
}

// sums the per-image head partials [n][625] -> grads: dw[ld], db[1], dpos[pts].  One warp per output: lanes take every
// 32nd image, then a fixed-order shuffle tree (deterministic).
__global__ void __launch_bounds__(256)
head_reduce_kernel(const float* __restrict__ partial, int n, int ld, int pts, float* __restrict__ dw, float* __restrict__ db,
                   float* __restrict__ dpos) {
    pdl_trigger();
    pdl_wait();
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= 256 + 1 + 368) return;
    float s = 0.f;
    for (int k = lane; k < n; k += 32) s += partial[(size_t)k * 625 + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane != 0) return;
    if (i < 256) { if (i < ld) dw[i] = s; }
    else if (i == 256) db[0] = s;
    else if (i - 257 < pts && dpos) dpos[i - 257] = s;
}

// Value head (value.py:23-28): 1x1 conv -> Flatten -> Dense(256, linear) -> Dense(1, tanh).  One CTA per image.
__global__ void __launch_bounds__(256)
value_head_kernel(const bf16* __restrict__ X, int ld, PixelLayout L, const float* __restrict__ w, const float* __restrict__ bias,
                  const float* __restrict__ W1, const float* __restrict__ b1, const float* __restrict__ w2,
                  const float* __restrict__ b2, int hidden, float* __restrict__ out) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s_v[368];
    __shared__ float s_w[256];
    __shared__ float s_red[8];
    const int img = blockIdx.x, S = L.size, pts = S * S;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = threadIdx.x; c < 256; c += 256) s_w[c] = c < ld ? w[c] : 0.f;
    __syncthreads();
    const int groups = ld / 8;
    for (int p = warp; p < pts; p += 8) {
        int x = p / S, y = p - x * S;
        const bf16* row = X + ((size_t)img * L.img_stride + (size_t)x * L.row_stride + y) * ld;
        float acc = 0.f;
        if (lane < groups) {
            uint4 v = *reinterpret_cast<const uint4*>(row + lane * 8);
            const bf16* vb = reinterpret_cast<const bf16*>(&v);
#pragma unroll
            for (int j = 0; j < 8; j++) acc += __bfloat162float(vb[j]) * s_w[lane * 8 + j];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) s_v[p] = acc + bias[0];
    }
    __syncthreads();
    float part = 0.f;
    for (int h = threadIdx.x; h < hidden; h += 256) {
        const float* wr = W1 + (size_t)h * pts;
        float a = b1[h];
        for (int p = 0; p < pts; p++) a += wr[p] * s_v[p];
        part += a * w2[h];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_red[warp] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = b2[0];
        for (int i = 0; i < 8; i++) t += s_red[i];
        out[img] = tanhf(t);
    }
}

// bias gradient: db[c] = sum_m G[m][c]; two deterministic stages (row slabs, then a warp per channel over the slabs)
__global__ void __launch_bounds__(256) colsum_stage1_kernel(const bf16* __restrict__ G, size_t rows, int ld, float* __restrict__ part) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s[256][9];
    const int groups = ld / 8, lanes = 256 / groups;
    const int g = threadIdx.x % groups, rl = threadIdx.x / groups;
    size_t per = (rows + gridDim.x - 1) / gridDim.x;
    size_t r0 = blockIdx.x * per, r1 = r0 + per < rows ? r0 + per : rows;
    float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (rl < lanes) {
        size_t r = r0 + rl;
        for (; r + 3 * (size_t)lanes < r1; r += 4 * (size_t)lanes) {  // four independent 16-byte loads in flight
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) v[u] = *reinterpret_cast<const uint4*>(G + (r + u * (size_t)lanes) * ld + g * 8);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bf16* vb = reinterpret_cast<const bf16*>(&v[u]);
#pragma unroll
                for (int j = 0; j < 8; j++) a[j] += __bfloat162float(vb[j]);
            }
        }
        for (; r < r1; r += lanes) {
            uint4 v = *reinterpret_cast<const uint4*>(G + r * ld + g * 8);
            const bf16* vb = reinterpret_cast<const bf16*>(&v);
#pragma unroll
            for (int j = 0; j < 8; j++) a[j] += __bfloat162float(vb[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < 8; j++) s[threadIdx.x][j] = a[j];
    __syncthreads();
    if (threadIdx.x < ld) {
        int gg = threadIdx.x / 8, j = threadIdx.x % 8;
        float t = 0.f;
        for (int k = 0; k < lanes; k++) t += s[k * groups + gg][j];
        part[(size_t)blockIdx.x * ld + threadIdx.x] = t;
    }
}
__global__ void __launch_bounds__(256) colsum_stage2_kernel(const float* __restrict__ part, int nblocks, int ld, int cout, float* __restrict__ db) {
    pdl_trigger();
    pdl_wait();
    const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (c >= cout) return;
    float t = 0.f;
    for (int b = lane; b < nblocks; b += 32) t += part[(size_t)b * ld + c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) db[c] = t;
}

// SGD with momentum (Keras 1.0 SGD.get_updates: v = momentum*v - lr*g; w += v) over EVERY parameter in one launch.
// Refreshes the bf16 operand copies: wf [tap][co][ci] (forward) and wb [tap][ci][co] (data gradient).
// One CTA per 32x32 (co, ci) tile of one tap of one layer: coalesced fp32 update along ci, the bf16 tile goes through
// shared memory so the transposed copy wb is written coalesced along co as well.  ldout and kin are multiples of 64.
// The CTA that owns a layer's first tile also updates its bias; the last few CTAs take the flat (head) parameters.
struct SgdLayer {
    float *w, *vel, *b, *bvel;
    const float *g, *gb;
    bf16 *wf, *wb;
    int taps, ldout, kin;
    int tile0;  // first global tile index of this layer
};
struct SgdFlat {
    float *w, *vel;
    const float* g;
    int count;
};
constexpr int kSgdMaxFlat = 4;
struct SgdFlats { SgdFlat f[kSgdMaxFlat]; int n; };

__global__ void __launch_bounds__(256)
sgd_all_kernel(const SgdLayer* __restrict__ layers, int n_layers, int total_tiles, const SgdFlats flats, float lr, float momentum, float gscale) {
    pdl_trigger();
    pdl_wait();
    // 64 x 64 (co, ci) tiles, four consecutive ci per thread: 16-byte loads of g / w (8 of them in flight per thread), 8-byte
    // bf16 stores, and the transposed copy through shared memory.  (Round 1: 32 x 32 tiles, scalar accesses: 26 us for 54 MB.)
    __shared__ __align__(8) bf16 tile[64][68];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // 16 ci quads x 16 rows
    for (int tidx = blockIdx.x; tidx < total_tiles + flats.n; tidx += gridDim.x) {
        if (tidx >= total_tiles) {  // a flat parameter vector
            const SgdFlat& f = flats.f[tidx - total_tiles];
            for (int i = threadIdx.x; i < f.count; i += blockDim.x) {
                float v = -lr * f.g[i] * gscale;
                if (momentum != 0.f) { v += momentum * f.vel[i]; f.vel[i] = v; }
                f.w[i] += v;
            }
            continue;
        }
        int l = 0;
        while (l + 1 < n_layers && layers[l + 1].tile0 <= tidx) l++;
        const SgdLayer L = layers[l];
        const int local = tidx - L.tile0;
        const int tiles_ci = L.kin / 64, tiles_co = L.ldout / 64;
        const int t = local / (tiles_co * tiles_ci);
        const int rem = local - t * tiles_co * tiles_ci;
        const int co0 = (rem / tiles_ci) * 64, ci0 = (rem % tiles_ci) * 64;
        float4 g[4], w[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const size_t i = ((size_t)t * L.ldout + co0 + ty + 16 * k) * L.kin + ci0 + 4 * tx;
            g[k] = *reinterpret_cast<const float4*>(L.g + i);
            w[k] = *reinterpret_cast<const float4*>(L.w + i);
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const size_t i = ((size_t)t * L.ldout + co0 + ty + 16 * k) * L.kin + ci0 + 4 * tx;
            float4 v = make_float4(-lr * g[k].x * gscale, -lr * g[k].y * gscale, -lr * g[k].z * gscale, -lr * g[k].w * gscale);
            if (momentum != 0.f) {  // plain SGD never touches the velocity
                const float4 o = *reinterpret_cast<const float4*>(L.vel + i);
                v.x += momentum * o.x; v.y += momentum * o.y; v.z += momentum * o.z; v.w += momentum * o.w;
                *reinterpret_cast<float4*>(L.vel + i) = v;
            }
            const float4 nw = make_float4(w[k].x + v.x, w[k].y + v.y, w[k].z + v.z, w[k].w + v.w);
            *reinterpret_cast<float4*>(L.w + i) = nw;
            const __nv_bfloat162 h01 = __floats2bfloat162_rn(nw.x, nw.y), h23 = __floats2bfloat162_rn(nw.z, nw.w);
            uint2 packed;
            packed.x = *reinterpret_cast<const uint32_t*>(&h01);
            packed.y = *reinterpret_cast<const uint32_t*>(&h23);
            *reinterpret_cast<uint2*>(L.wf + i) = packed;
            *reinterpret_cast<uint2*>(&tile[ty + 16 * k][4 * tx]) = packed;
        }
        __syncthreads();
        if (L.wb) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int ci = ty + 16 * k;   // four consecutive co of one ci per thread
                const __nv_bfloat162 a = __halves2bfloat162(tile[4 * tx][ci], tile[4 * tx + 1][ci]);
                const __nv_bfloat162 b = __halves2bfloat162(tile[4 * tx + 2][ci], tile[4 * tx + 3][ci]);
                uint2 packed;
                packed.x = *reinterpret_cast<const uint32_t*>(&a);
                packed.y = *reinterpret_cast<const uint32_t*>(&b);
                *reinterpret_cast<uint2*>(L.wb + ((size_t)t * L.kin + ci0 + ci) * L.ldout + co0 + 4 * tx) = packed;
            }
        }
        __syncthreads();
        if (local == 0) {
            for (int j = threadIdx.x; j < L.ldout; j += blockDim.x) {
                float v = -lr * L.gb[j] * gscale;
                if (momentum != 0.f) { v += momentum * L.bvel[j]; L.bvel[j] = v; }
                L.b[j] += v;
            }
        }
    }
}

__global__ void unpack_rows_kernel(const bf16* __restrict__ rows, int n, int C, int S, PixelLayout L, int ld, float* __restrict__ out) {
    size_t total = (size_t)n * C * S * S;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int p = (int)(i % (S * S));
        size_t r = i / (S * S);
        int c = (int)(r % C);
        int img = (int)(r / C);
        int x = p / S, y = p - x * S;
        out[i] = __bfloat162float(rows[((size_t)img * L.img_stride + (size_t)x * L.row_stride + y) * ld + c]);
    }
}

// ------------------------------------------------------------------------------------ net object

struct ConvLayer {
    int cin, cout, width, taps, pad;
    int kin;    // input row length (cin padded to 64)
    int ldout;  // output row length (cout padded to 64)
    int nout;   // UMMA N forward (cout padded to 16)
    int nin;    // UMMA N of the data gradient (cin padded to 16)
    PixelLayout lin, lout;
    size_t wcount;  // taps*ldout*kin
    float *w, *b, *vw, *vb;  // fp32 master weights [tap][ldout][kin], bias [ldout], velocities
    float *gw, *gb;          // slices of the flat gradient buffer
    bf16 *wf, *wb;           // bf16 operand copies (wb null for layer 0)
    bf16* act;               // output activations [max_batch*lout.img_stride][ldout]
    uint32_t* mask;          // ReLU bitmask of `act`, kConvMaskWords words per row (written by the forward conv, read by the data gradient above)
    int off_fwd[kConvMaxTaps], off_bwd[kConvMaxTaps];
    CUtensorMap tm_in, tm_wf, tm_g, tm_wb;      // forward A / B, dgrad A / B (full-size tensors)
    CUtensorMap tm_wf_half, tm_wb_half;         // B with a box of N/2 rows, for the CTA-pair kernel
    CUtensorMap tm_in_win, tm_g_win;            // A with a box of win_rows rows (activation window kernel)
    bool win_fwd, win_bwd;
    bool pair_fwd, pair_bwd;
    WgradMaps tm_wg;                            // wgrad operands, exact batch rows (rebuilt when n changes)
    int wgrad_splits;
    bool wgrad_chain;                           // can take part in a chained weight-gradient launch
    unsigned int wgrad_chain_count;             // chained launches since its counters were zeroed
    unsigned int wgrad_done_target;             // value of its "done" counter once the latest chained launch has finished it
};

}  // namespace ag

using namespace ag;

constexpr int kWgradCounterStride = 64;   // >= tiles of any chained layer

struct ag_net {
    int board, in_planes, n_conv, head_kind, max_batch;
    std::vector<ConvLayer> L;
    PixelLayout lay_in, lay_mid;
    bf16* x0 = nullptr;        // packed input rows [max_batch*lay_in.img_stride][kin0]
    bf16* x0_grad = nullptr;   // G of layer 0 in the input layout (only when layer 0's layout differs)
    std::vector<bf16*> gbuf;   // per-layer pre-activation gradients in lay_mid ([0] unused when x0_grad holds layer 0's)
    // layer chains (conv.cu, conv_chain_kernel): the whole forward stack / data-gradient chain in one launch each
    bool chain_ok = false;      // the layers can run as chains
    bool chain_fwd = false, chain_bwd = false;  // ... and the forward stack / the data gradients do (AG_CONV_CHAIN: 1 both, 2 forward only, 3 backward only)
    int chain_n = -1;          // batch the device tables were built for
    std::vector<ChainParams> chain_fwd_p, chain_bwd_p;   // launches of <= kChainMaxLayers layers each
    unsigned int *chain_flags_fwd = nullptr, *chain_flags_bwd = nullptr;
    unsigned int chain_epoch_fwd = 0, chain_epoch_bwd = 0;
    int chain_win_rows = 0, chain_nmax_fwd = 0, chain_nmax_bwd = 0;
    bool dgrad_chain_done = false;  // the data-gradient chain of the current forward_loss has been queued
    cudaEvent_t ev_first = nullptr;     // recorded after the first layer's weight-gradient kernels when they are queued ahead of the chain
    bool first_layer_early = false;     // ... in the latest backward pass
    bool wgrad_chain_refused = false;   // a chained weight-gradient launch was refused (grid not co-resident): stay on per-layer kernels
    int chain_fallbacks = 0;        // chained launches the device refused, each answered by the per-layer kernels
    // head
    float *hw = nullptr, *hb = nullptr, *hpos = nullptr;       // 1x1 conv w[ld], b[1], position bias[pts] (optional)
    float *vhw = nullptr, *vhb = nullptr, *vhpos = nullptr;    // velocities
    float *g_hw = nullptr, *g_hb = nullptr, *g_hpos = nullptr;  // gradient slices
    int use_pos_bias = 0;
    float *d1w = nullptr, *d1b = nullptr, *d2w = nullptr, *d2b = nullptr;  // value head dense layers
    int hidden = 256;
    float* head_partial = nullptr;  // [max_batch][625]
    float* colsum_part = nullptr;   // [colsum_blocks][256]
    float* wgrad_scratch = nullptr;
    float* wgrad_scratch2 = nullptr;          // second partial buffer: chained weight gradients alternate between the two
    unsigned int* wgrad_counters = nullptr;   // [n_conv][kWgradCounterStride] split-arrival counters of the chained launches
    size_t wgrad_scratch_floats = 0;
    float* grads = nullptr;  // flat: per conv layer gw then gb; then head w, b, pos
    size_t grad_count = 0;
    std::vector<long long> grad_offsets;  // start of each conv layer's slice, then head, then end
    int wgrad_n = -1;
    int sm_budget = 0;   // SMs the weight-gradient launches may fill (0 = all)
    int loss_kind = 0;   // 0 categorical cross-entropy (SL trainer), 1 Keras binary cross-entropy on the softmax output (RL trainer)
    int colsum_blocks = 0;
    long long iterations = 0;
    void* sgd_table = nullptr;      // device copy of the SGD kernel's per-layer table
    int sgd_tiles = 0;
};

namespace ag {

template <class T>
static int dev_alloc(T** p, size_t count, bool zero = true) {
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T) + 256);
    if (e != cudaSuccess) { set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e)); return AG_ENOMEM; }
    if (zero) { e = cudaMemset(*p, 0, count * sizeof(T) + 256); if (e != cudaSuccess) return cuda_fail(e, "cudaMemset"); }
    return AG_OK;
}

static void tap_offsets(int width, const PixelLayout& L, int* fwd, int* bwd) {
    int pad = width / 2, t = 0;
    for (int kx = 0; kx < width; kx++)
        for (int ky = 0; ky < width; ky++, t++) {
            fwd[t] = (kx - pad) * L.row_stride + (ky - pad);  // cross-correlation tap (kx, ky)
            bwd[t] = -fwd[t];
        }
}

static int grid_for(size_t work, int threads) {
    size_t g = (work + threads - 1) / threads;
    size_t cap = (size_t)sm_count() * 16;
    return (int)(g < cap ? (g ? g : 1) : cap);
}

static void fwd_conv_args(ag_net* net, int l, int n, ConvArgs* a) {
    ConvLayer& c = net->L[l];
    memset(a, 0, sizeof(*a));
    a->m_total = n * c.lin.img_stride;
    a->n = c.nout; a->taps = c.taps; a->kchunks = c.kin / 64;
    memcpy(a->off, c.off_fwd, sizeof(int) * c.taps);
    a->in = c.lin; a->dst = c.lout;
    a->bias = c.b; a->mask_out = c.mask; a->out = c.act; a->out_ld = c.ldout; a->b_tap_rows = c.ldout; a->relu = 1;
}

// data gradient of layer `l` (> 0) into the layer below, gated by that layer's ReLU mask
static void bwd_conv_args(ag_net* net, int l, int n, ConvArgs* a) {
    ConvLayer& c = net->L[l];
    ConvLayer& below = net->L[l - 1];
    memset(a, 0, sizeof(*a));
    a->m_total = n * c.lout.img_stride;
    a->n = c.nin; a->taps = c.taps; a->kchunks = c.ldout / 64;
    memcpy(a->off, c.off_bwd, sizeof(int) * c.taps);
    a->in = c.lout;
    a->gate_mask = below.mask; a->gate_layout = below.lout;
    a->out_ld = below.ldout; a->b_tap_rows = c.kin;
    if (l - 1 == 0 && net->x0_grad) {
        // the first layer's G lives in the INPUT layout so its weight gradient can pair it with x0; the gate
        // (layer-0 ReLU mask) is still read in the mid layout
        a->dst = net->lay_in; a->out = net->x0_grad;
    } else {
        a->dst = c.lout; a->out = net->gbuf[l - 1];
    }
}

// half-tiles per image when a layer's tiles are aligned to images (0: plain 128-row tiling), see ChainLayer::tiles_img
static int chain_tiles_img(const PixelLayout& L) {
    static int on = -1;
    if (on < 0) { const char* e = getenv("AG_CHAIN_IMG_TILES"); on = (e && e[0] == '0') ? 0 : 1; }
    const int k = (L.size * L.row_stride + 127) / 128;   // rows of an image that produce output
    return (on && k * 128 < L.img_stride) ? k : 0;
}

static void chain_set_tiles(ChainLayer& e, int n_images) {
    e.tiles_img = chain_tiles_img(e.a.in);
    const int halves = e.tiles_img ? n_images * e.tiles_img : (e.a.m_total + 127) / 128;
    e.n_pairs = (halves + 1) / 2;
}

// kernel-parameter tables of the two layer chains for batch n, split into launches of <= kChainMaxLayers layers
static void chain_append(std::vector<ChainParams>& out, const ChainLayer& e_in, int* flag0) {
    if (out.empty() || out.back().n_layers == kChainMaxLayers) {
        out.emplace_back();
        memset(&out.back(), 0, sizeof(ChainParams));
    }
    ChainParams& P = out.back();
    ChainLayer e = e_in;
    e.tile0 = P.total_tiles;
    e.flag0 = *flag0;
    if (P.n_layers == 0) e.dep = 0;  // the launch before this one has finished
    P.layers[P.n_layers++] = e;
    P.total_tiles += e.n_pairs;
    *flag0 += 2 * e.n_pairs;
}

static int net_build_chains(ag_net* net, int n) {
    if (!net->chain_ok || net->chain_n == n) return AG_OK;
    const int nc = net->n_conv;
    net->chain_fwd_p.clear();
    net->chain_bwd_p.clear();
    int flag0 = 0;
    for (int l = 0; l < nc; l++) {
        ChainLayer e;
        memset(&e, 0, sizeof(e));
        ConvLayer& c = net->L[l];
        e.tmap_aw = c.tm_in_win; e.tmap_bh = c.tm_wf_half;
        fwd_conv_args(net, l, n, &e.a);
        e.a.ksteps_last = (c.cin % 64) ? ((c.cin % 64) + 15) / 16 : 0;   // input channels past cin are zero padding
        e.a.win_rows = conv_window_rows(e.a.off, e.a.taps); e.a.halo = (e.a.win_rows - 128) / 2;
        chain_set_tiles(e, n);
        if (l > 0) {
            ConvLayer& pv = net->L[l - 1];
            e.dep = (pv.lin.img_stride == pv.lout.img_stride && pv.lin.row_stride == pv.lout.row_stride) ? 1 : 2;
            e.prev_in_img_stride = pv.lin.img_stride; e.prev_rows = n * pv.lin.img_stride;
            e.prev_tiles_img = chain_tiles_img(pv.lin);
        }
        chain_append(net->chain_fwd_p, e, &flag0);
    }
    flag0 = 0;
    for (int k = 0; k + 1 < nc; k++) {
        const int l = nc - 1 - k;
        ChainLayer e;
        memset(&e, 0, sizeof(e));
        ConvLayer& c = net->L[l];
        e.tmap_aw = c.tm_g_win; e.tmap_bh = c.tm_wb_half;
        bwd_conv_args(net, l, n, &e.a);
        e.a.win_rows = conv_window_rows(e.a.off, e.a.taps); e.a.halo = (e.a.win_rows - 128) / 2;
        chain_set_tiles(e, n);
        if (k > 0) {  // G rows all live in lay_mid
            e.dep = 1; e.prev_in_img_stride = c.lout.img_stride; e.prev_rows = n * c.lout.img_stride;
            e.prev_tiles_img = chain_tiles_img(c.lout);
        }
        chain_append(net->chain_bwd_p, e, &flag0);
    }
    net->chain_n = n;
    return AG_OK;
}

// K blocks per image for the weight gradient when they are aligned to images (0: linear), see WgradArgs::kb_img
static int wgrad_kb_img(const PixelLayout& L, int krows) {
    static int on = -1;
    if (on < 0) { const char* e = getenv("AG_WGRAD_IMG_BLOCKS"); on = (e && e[0] == '0') ? 0 : 1; }
    const int k = (L.size * L.row_stride + krows - 1) / krows;
    return (on && k * krows < L.img_stride) ? k : 0;
}

static int net_build_wgrad_maps(ag_net* net, int n) {
    if (net->wgrad_n == n) return AG_OK;
    for (int l = 0; l < net->n_conv; l++) {
        ConvLayer& c = net->L[l];
        const bf16* xin = l == 0 ? net->x0 : net->L[l - 1].act;
        const bf16* g = (l == 0 && net->x0_grad) ? net->x0_grad : net->gbuf[l];
        uint64_t rows = (uint64_t)n * c.lin.img_stride;
        int rc = wgrad_make_maps(&c.tm_wg, g, xin, rows, c.ldout, c.kin, c.taps, c.off_fwd);
        if (rc) return rc;
        c.wgrad_splits = wgrad_plan_splits(c.taps, c.ldout, c.kin, (int)rows, net->sm_budget);
        size_t need = (size_t)c.wgrad_splits * wgrad_partial_floats(c.taps, c.ldout, c.kin);
        if (need > net->wgrad_scratch_floats) { set_error("wgrad scratch too small"); return AG_ESTATE; }
        WgradArgs w;
        memset(&w, 0, sizeof(w));
        w.m_total = (int)rows; w.taps = c.taps; w.ldout = c.ldout; w.kin = c.kin; w.splits = c.wgrad_splits;
        w.kb_img = wgrad_kb_img(c.lin, c.tm_wg.krows); w.img_stride = c.lin.img_stride;
        const int wg_tiles = (c.taps * (c.ldout / 64) + 3) / 4;   // the chained kernel keeps whole-accumulator partials: 256 rows x kin per tile
        c.wgrad_chain = !net->wgrad_chain_refused && wgrad_chain_ok(w, c.tm_wg, c.kin + 16 <= 256) && wg_tiles < kWgradCounterStride &&
                        (size_t)c.wgrad_splits * wg_tiles * 256 * c.kin <= net->wgrad_scratch_floats;
    }
    // the arrival counters count in units of this batch's split factors: start over.  (Batch-size changes are rare; the device
    // is drained first so that no weight-gradient launch of the old batch size, on whatever stream, still uses them.)
    AG_CUDA(cudaDeviceSynchronize());
    AG_CUDA(cudaMemset(net->wgrad_counters, 0, sizeof(unsigned int) * net->n_conv * kWgradCounterStride));
    for (auto& c : net->L) { c.wgrad_chain_count = 0; c.wgrad_done_target = 0; }
    net->wgrad_n = n;
    return AG_OK;
}

}  // namespace ag

extern "C" {

ag_net* ag_net_create(int board, int in_planes, int n_conv, const int* h_filters, const int* h_widths, int head_kind, int max_batch) {
    if (board < 2 || board > 19 || in_planes < 1 || in_planes > 256 || n_conv < 1 || n_conv > 64 || max_batch < 1 ||
        (head_kind != 0 && head_kind != 1)) {
        set_error("ag_net_create: bad argument");
        return nullptr;
    }
    int midpad = 0;
    for (int l = 0; l < n_conv; l++) {
        if (h_widths[l] < 1 || h_widths[l] > 5 || h_widths[l] % 2 == 0) { set_error("ag_net_create: filter width %d must be odd and <= 5", h_widths[l]); return nullptr; }
        if (h_filters[l] < 1 || h_filters[l] > 256) { set_error("ag_net_create: %d filters unsupported (1..256)", h_filters[l]); return nullptr; }
        if (l > 0 && h_widths[l] / 2 > midpad) midpad = h_widths[l] / 2;
        if (h_filters[l] != h_filters[0]) { set_error("ag_net_create: all conv layers must have the same filter count (policy.py:97-103 filters_per_layer)"); return nullptr; }
    }
    if (midpad == 0) midpad = 0;
    ag_net* net = new ag_net();
    net->board = board; net->in_planes = in_planes; net->n_conv = n_conv; net->head_kind = head_kind; net->max_batch = max_batch;
    int pad0 = h_widths[0] / 2;
    net->lay_in = PixelLayout{board, board + pad0, (board + pad0) * (board + pad0)};
    net->lay_mid = PixelLayout{board, board + midpad, (board + midpad) * (board + midpad)};
    if (pad0 == midpad) net->lay_in = net->lay_mid;
    const bool same_layout = net->lay_in.row_stride == net->lay_mid.row_stride;
    net->L.resize(n_conv);
    int rc = AG_OK;
    size_t gcount = 0, scratch = 0;
    for (int l = 0; l < n_conv && !rc; l++) {
        ConvLayer& c = net->L[l];
        c.cin = l == 0 ? in_planes : h_filters[l - 1];
        c.cout = h_filters[l];
        c.width = h_widths[l]; c.taps = c.width * c.width; c.pad = c.width / 2;
        c.kin = round_up(c.cin, 64); c.ldout = round_up(c.cout, 64);
        c.nout = round_up(c.cout, 16); c.nin = round_up(c.cin, 16);
        c.lin = l == 0 ? net->lay_in : net->lay_mid;
        c.lout = net->lay_mid;
        c.wcount = (size_t)c.taps * c.ldout * c.kin;
        tap_offsets(c.width, c.lin, c.off_fwd, c.off_bwd);
        net->grad_offsets.push_back((long long)gcount);
        gcount += c.wcount + c.ldout;
        size_t rows_in = (size_t)max_batch * c.lin.img_stride;
        int splits = wgrad_plan_splits(c.taps, c.ldout, c.kin, (int)rows_in, 0);
        const size_t partials = (size_t)kWgradMaxSplits * wgrad_partial_floats(c.taps, c.ldout, c.kin);
        if (partials > scratch) scratch = partials;
        (void)splits;
    }
    net->grad_offsets.push_back((long long)gcount);  // head slice
    const int ldh = net->L[n_conv - 1].ldout, pts = board * board;
    gcount += 256 + 1 + 368;
    gcount = (gcount + 3) & ~(size_t)3;   // whole 16-byte groups (the head's 625 values + 3 unused zeros): the exchange moves float4
    net->grad_offsets.push_back((long long)gcount);
    net->grad_count = gcount;
    rc = dev_alloc(&net->grads, gcount);
    net->wgrad_scratch_floats = scratch;
    if (!rc) rc = dev_alloc(&net->wgrad_scratch, scratch + (size_t)kWgradMaxSplits * 256, false);  // + bias-column partials
    if (!rc) rc = dev_alloc(&net->wgrad_scratch2, scratch + (size_t)kWgradMaxSplits * 256, false);
    if (!rc) rc = dev_alloc(&net->wgrad_counters, (size_t)n_conv * kWgradCounterStride);
    if (!rc) rc = dev_alloc(&net->x0, (size_t)max_batch * net->lay_in.img_stride * net->L[0].kin);
    if (!rc && !same_layout) rc = dev_alloc(&net->x0_grad, (size_t)max_batch * net->lay_in.img_stride * net->L[0].ldout);
    net->gbuf.assign(n_conv, nullptr);
    for (int l = 0; l < n_conv && !rc; l++)
        if (!(l == 0 && net->x0_grad)) rc = dev_alloc(&net->gbuf[l], (size_t)max_batch * net->lay_mid.img_stride * net->L[l].ldout);
    for (int l = 0; l < n_conv && !rc; l++) {
        ConvLayer& c = net->L[l];
        rc = dev_alloc(&c.w, c.wcount);
        if (!rc) rc = dev_alloc(&c.b, (size_t)c.ldout);
        if (!rc) rc = dev_alloc(&c.vw, c.wcount);
        if (!rc) rc = dev_alloc(&c.vb, (size_t)c.ldout);
        if (!rc) rc = dev_alloc(&c.wf, c.wcount);
        c.wb = nullptr;
        if (!rc && l > 0) rc = dev_alloc(&c.wb, c.wcount);
        if (!rc) rc = dev_alloc(&c.act, (size_t)max_batch * c.lout.img_stride * c.ldout);
        if (!rc) rc = dev_alloc(&c.mask, (size_t)max_batch * c.lout.img_stride * kConvMaskWords);
        c.gw = net->grads + net->grad_offsets[l];
        c.gb = c.gw + c.wcount;
        if (rc) break;
        const bf16* xin = l == 0 ? net->x0 : net->L[l - 1].act;
        rc = make_tmap_2d(&c.tm_in, xin, (uint64_t)max_batch * c.lin.img_stride, c.kin, 128);
        if (!rc) rc = make_tmap_2d(&c.tm_wf, c.wf, (uint64_t)c.taps * c.ldout, c.kin, c.nout);
        c.pair_fwd = c.nout % 32 == 0;
        c.pair_bwd = l > 0 && c.nin % 32 == 0;
        if (!rc && c.pair_fwd) rc = make_tmap_2d(&c.tm_wf_half, c.wf, (uint64_t)c.taps * c.ldout, c.kin, c.nout / 2);
        const int win_rows = conv_window_rows(c.off_fwd, c.taps);
        c.win_fwd = c.pair_fwd && win_rows > 0;
        c.win_bwd = c.pair_bwd && win_rows > 0;
        if (!rc && c.win_fwd) rc = make_tmap_2d(&c.tm_in_win, xin, (uint64_t)max_batch * c.lin.img_stride, c.kin, win_rows);
        if (!rc && c.win_bwd) rc = make_tmap_2d(&c.tm_g_win, net->gbuf[l], (uint64_t)max_batch * c.lout.img_stride, c.ldout, win_rows);
        if (!rc && l > 0) {
            // data gradient: A = G of this layer (lay_mid rows, ldout wide), B = wb [tap][ci][co]
            rc = make_tmap_2d(&c.tm_g, net->gbuf[l], (uint64_t)max_batch * c.lout.img_stride, c.ldout, 128);
            if (!rc) rc = make_tmap_2d(&c.tm_wb, c.wb, (uint64_t)c.taps * c.kin, c.ldout, c.nin);
            if (!rc && c.pair_bwd) rc = make_tmap_2d(&c.tm_wb_half, c.wb, (uint64_t)c.taps * c.kin, c.ldout, c.nin / 2);
        }
    }
    if (!rc) {  // layer chains: every layer must be able to run the pair + window kernel
        bool ok = conv_chain_enabled();
        int wmax = 0, nf = 0, nb = 0;
        size_t flags = 0;
        for (int l = 0; l < n_conv; l++) {
            ConvLayer& c = net->L[l];
            ok = ok && c.win_fwd && (l == 0 || c.win_bwd);
            int w = conv_window_rows(c.off_fwd, c.taps);
            if (w > wmax) wmax = w;
            if (c.nout > nf) nf = c.nout;
            if (l > 0 && c.nin > nb) nb = c.nin;
            size_t rows = (size_t)max_batch * (c.lin.img_stride > c.lout.img_stride ? c.lin.img_stride : c.lout.img_stride);
            flags += 2 * ((rows + 255) / 256);
        }
        net->chain_ok = ok;
        { const char* e = getenv("AG_CONV_CHAIN"); int m = e ? atoi(e) : 1; net->chain_fwd = ok && (m == 1 || m == 2); net->chain_bwd = ok && (m == 1 || m == 3); }
        net->chain_win_rows = wmax; net->chain_nmax_fwd = nf; net->chain_nmax_bwd = nb;
        if (ok) {
            rc = dev_alloc(&net->chain_flags_fwd, flags);
            if (!rc) rc = dev_alloc(&net->chain_flags_bwd, flags);
        }
    }
    if (!rc) {  // the SGD kernel's layer table
        std::vector<SgdLayer> tab(n_conv);
        int tile0 = 0;
        for (int l = 0; l < n_conv; l++) {
            ConvLayer& c = net->L[l];
            tab[l] = SgdLayer{c.w, c.vw, c.b, c.vb, c.gw, c.gb, c.wf, c.wb, c.taps, c.ldout, c.kin, tile0};
            tile0 += c.taps * (c.ldout / 64) * (c.kin / 64);   // 64 x 64 tiles (ldout, kin are multiples of 64)
        }
        net->sgd_tiles = tile0;
        { unsigned char* t = nullptr; rc = dev_alloc(&t, sizeof(SgdLayer) * n_conv); net->sgd_table = t; }
        if (!rc && cudaMemcpy(net->sgd_table, tab.data(), sizeof(SgdLayer) * n_conv, cudaMemcpyHostToDevice) != cudaSuccess) rc = AG_ECUDA;
    }
    if (!rc) rc = dev_alloc(&net->hw, 256);
    if (!rc) rc = dev_alloc(&net->hb, 4);
    if (!rc) rc = dev_alloc(&net->hpos, 368);
    if (!rc) rc = dev_alloc(&net->vhw, 256);
    if (!rc) rc = dev_alloc(&net->vhb, 4);
    if (!rc) rc = dev_alloc(&net->vhpos, 368);
    if (!rc) rc = dev_alloc(&net->head_partial, (size_t)max_batch * 625);
    net->colsum_blocks = sm_count() * 8;
    if (!rc) rc = dev_alloc(&net->colsum_part, (size_t)net->colsum_blocks * 256);
    if (!rc && head_kind == 1) {
        rc = dev_alloc(&net->d1w, (size_t)net->hidden * pts);
        if (!rc) rc = dev_alloc(&net->d1b, (size_t)net->hidden);
        if (!rc) rc = dev_alloc(&net->d2w, (size_t)net->hidden);
        if (!rc) rc = dev_alloc(&net->d2b, 4);
    }
    if (!rc) {
        net->g_hw = net->grads + net->grad_offsets[n_conv];
        net->g_hb = net->g_hw + 256;
        net->g_hpos = net->g_hb + 1;
        (void)ldh;
        if (cudaDeviceSynchronize() != cudaSuccess) rc = AG_ECUDA;
    }
    if (rc) { ag_net_destroy(net); return nullptr; }
    return net;
}

void ag_net_destroy(ag_net* net) {
    if (net && net->ev_first) cudaEventDestroy(net->ev_first);
    if (!net) return;
    for (auto& c : net->L) { cudaFree(c.w); cudaFree(c.b); cudaFree(c.vw); cudaFree(c.vb); cudaFree(c.wf); cudaFree(c.wb); cudaFree(c.act); cudaFree(c.mask); }
    float* f[] = {net->hw, net->hb, net->hpos, net->vhw, net->vhb, net->vhpos, net->d1w, net->d1b, net->d2w, net->d2b,
                  net->head_partial, net->colsum_part, net->wgrad_scratch, net->wgrad_scratch2, net->grads};
    cudaFree(net->wgrad_counters);
    for (float* p : f) cudaFree(p);
    cudaFree(net->sgd_table); cudaFree(net->x0); cudaFree(net->x0_grad);
    for (bf16* g : net->gbuf) cudaFree(g);
    cudaFree(net->chain_flags_fwd); cudaFree(net->chain_flags_bwd);
    delete net;
}

// layer geometry for the Python side: field 0 cin, 1 cout, 2 width, 3 kin, 4 ldout
int ag_net_layer_info(const ag_net* net, int layer, int field) {
    if (layer < 0 || layer >= net->n_conv) return -1;
    const ConvLayer& c = net->L[layer];
    switch (field) {
        case 0: return c.cin; case 1: return c.cout; case 2: return c.width; case 3: return c.kin; case 4: return c.ldout;
        case 5: return net->chain_fwd ? 1 : 0;                       // forward runs inside the chained launch
        case 6: return (net->chain_bwd && layer > 0) ? 1 : 0;        // its data gradient does
        case 7: return (c.wgrad_chain && net->wgrad_n > 0) ? 1 : 0;  // its weight gradient does (known after a forward_loss)
        case 8: return net->chain_fallbacks;                         // chained launches refused so far (whole net)
        case 9: return c.wgrad_splits;                               // split-K factor planned for its weight gradient
    }
    return -1;
}

// Conv weights as cross-correlation filters in Keras 'th' order OIHW: w[co][ci][kx][ky], b[co]
// (policy.py:119-150 layers).  A Theano-trained file stores true-convolution filters: flip both
// spatial axes before calling (the Python layer's flip_kernels switch).  Synchronous.
int ag_net_set_conv(ag_net* net, int layer, const float* h_w, const float* h_b) {
    if (layer < 0 || layer >= net->n_conv) { set_error("bad layer %d", layer); return AG_EINVAL; }
    ConvLayer& c = net->L[layer];
    std::vector<float> w(c.wcount, 0.f), b(c.ldout, 0.f);
    std::vector<bf16> wf(c.wcount), wb(c.wcount);
    for (int co = 0; co < c.cout; co++) {
        b[co] = h_b ? h_b[co] : 0.f;
        for (int ci = 0; ci < c.cin; ci++)
            for (int t = 0; t < c.taps; t++)
                w[((size_t)t * c.ldout + co) * c.kin + ci] = h_w[((size_t)co * c.cin + ci) * c.taps + t];
    }
    for (size_t i = 0; i < c.wcount; i++) {
        wf[i] = __float2bfloat16(w[i]);
        int ci = (int)(i % c.kin); size_t r = i / c.kin; int co = (int)(r % c.ldout); int t = (int)(r / c.ldout);
        wb[((size_t)t * c.kin + ci) * c.ldout + co] = wf[i];
    }
    AG_CUDA(cudaMemcpy(c.w, w.data(), c.wcount * 4, cudaMemcpyHostToDevice));
    AG_CUDA(cudaMemcpy(c.b, b.data(), (size_t)c.ldout * 4, cudaMemcpyHostToDevice));
    AG_CUDA(cudaMemcpy(c.wf, wf.data(), c.wcount * 2, cudaMemcpyHostToDevice));
    if (c.wb) AG_CUDA(cudaMemcpy(c.wb, wb.data(), c.wcount * 2, cudaMemcpyHostToDevice));
    AG_CUDA(cudaMemset(c.vw, 0, c.wcount * 4));
    AG_CUDA(cudaMemset(c.vb, 0, (size_t)c.ldout * 4));
    return AG_OK;
}
int ag_net_get_conv(ag_net* net, int layer, float* h_w, float* h_b) {
    if (layer < 0 || layer >= net->n_conv) { set_error("bad layer %d", layer); return AG_EINVAL; }
    ConvLayer& c = net->L[layer];
    std::vector<float> w(c.wcount), b(c.ldout);
    AG_CUDA(cudaMemcpy(w.data(), c.w, c.wcount * 4, cudaMemcpyDeviceToHost));
    AG_CUDA(cudaMemcpy(b.data(), c.b, (size_t)c.ldout * 4, cudaMemcpyDeviceToHost));
    for (int co = 0; co < c.cout; co++) {
        if (h_b) h_b[co] = b[co];
        for (int ci = 0; ci < c.cin; ci++)
            for (int t = 0; t < c.taps; t++)
                h_w[((size_t)co * c.cin + ci) * c.taps + t] = w[((size_t)t * c.ldout + co) * c.kin + ci];
    }
    return AG_OK;
}
// head parameters.  which: 0 = 1x1 conv (w[C_last], b[1]); 1 = position bias (w[board*board], enables it);
// 2 = value Dense(256) (w[256][board*board], b[256]); 3 = value Dense(1) (w[256], b[1]).  Synchronous.
int ag_net_set_head(ag_net* net, int which, const float* h_w, const float* h_b) {
    const int pts = net->board * net->board, cl = net->L[net->n_conv - 1].cout;
    switch (which) {
        case 0:
            AG_CUDA(cudaMemset(net->hw, 0, 256 * 4));
            AG_CUDA(cudaMemcpy(net->hw, h_w, (size_t)cl * 4, cudaMemcpyHostToDevice));
            if (h_b) AG_CUDA(cudaMemcpy(net->hb, h_b, 4, cudaMemcpyHostToDevice));
            AG_CUDA(cudaMemset(net->vhw, 0, 256 * 4)); AG_CUDA(cudaMemset(net->vhb, 0, 4));
            return AG_OK;
        case 1:
            AG_CUDA(cudaMemcpy(net->hpos, h_w, (size_t)pts * 4, cudaMemcpyHostToDevice));
            AG_CUDA(cudaMemset(net->vhpos, 0, 368 * 4));
            net->use_pos_bias = 1;
            return AG_OK;
        case 2:
            if (net->head_kind != 1) break;
            AG_CUDA(cudaMemcpy(net->d1w, h_w, (size_t)net->hidden * pts * 4, cudaMemcpyHostToDevice));
            if (h_b) AG_CUDA(cudaMemcpy(net->d1b, h_b, (size_t)net->hidden * 4, cudaMemcpyHostToDevice));
            return AG_OK;
        case 3:
            if (net->head_kind != 1) break;
            AG_CUDA(cudaMemcpy(net->d2w, h_w, (size_t)net->hidden * 4, cudaMemcpyHostToDevice));
            if (h_b) AG_CUDA(cudaMemcpy(net->d2b, h_b, 4, cudaMemcpyHostToDevice));
            return AG_OK;
    }
    set_error("ag_net_set_head: bad selector %d for this head", which);
    return AG_EINVAL;
}
int ag_net_get_head(ag_net* net, int which, float* h_w, float* h_b) {
    const int pts = net->board * net->board, cl = net->L[net->n_conv - 1].cout;
    switch (which) {
        case 0:
            AG_CUDA(cudaMemcpy(h_w, net->hw, (size_t)cl * 4, cudaMemcpyDeviceToHost));
            if (h_b) AG_CUDA(cudaMemcpy(h_b, net->hb, 4, cudaMemcpyDeviceToHost));
            return AG_OK;
        case 1:
            AG_CUDA(cudaMemcpy(h_w, net->hpos, (size_t)pts * 4, cudaMemcpyDeviceToHost));
            return AG_OK;
        case 2:
            if (net->head_kind != 1) break;
            AG_CUDA(cudaMemcpy(h_w, net->d1w, (size_t)net->hidden * pts * 4, cudaMemcpyDeviceToHost));
            if (h_b) AG_CUDA(cudaMemcpy(h_b, net->d1b, (size_t)net->hidden * 4, cudaMemcpyDeviceToHost));
            return AG_OK;
        case 3:
            if (net->head_kind != 1) break;
            AG_CUDA(cudaMemcpy(h_w, net->d2w, (size_t)net->hidden * 4, cudaMemcpyDeviceToHost));
            if (h_b) AG_CUDA(cudaMemcpy(h_b, net->d2b, 4, cudaMemcpyDeviceToHost));
            return AG_OK;
    }
    set_error("ag_net_get_head: bad selector %d for this head", which);
    return AG_EINVAL;
}

static int net_conv_stack(ag_net* net, const uint8_t* d_planes, int n, cudaStream_t st) {
    const int S = net->board;
    ConvLayer& c0 = net->L[0];
    size_t work = (size_t)n * S * S * (c0.kin / 8);
    launch_k(pack_input_kernel, dim3(grid_for(work, 256)), dim3(256), 0, st, d_planes, n, net->in_planes, S, net->lay_in, c0.kin, net->x0);
    AG_LAUNCH_CHECK("pack_input_kernel");
    if (net->chain_fwd) {
        int rc = net_build_chains(net, n);
        if (rc) return rc;
        ++net->chain_epoch_fwd;
        for (const ChainParams& P : net->chain_fwd_p) {
            rc = conv_chain_launch(P, net->chain_flags_fwd, net->chain_epoch_fwd, net->chain_win_rows, net->chain_nmax_fwd, st);
            if (rc == AG_ESTATE) break;   // the device cannot hold the whole grid at once: per-layer kernels from here on
            if (rc) return rc;
        }
        if (rc == AG_OK) return AG_OK;
        net->chain_fwd = false;
        net->chain_fallbacks++;
    }
    for (int l = 0; l < net->n_conv; l++) {
        ConvLayer& c = net->L[l];
        ConvArgs a;
        fwd_conv_args(net, l, n, &a);
        int rc = conv_gemm_launch(c.tm_in, c.tm_wf, c.pair_fwd ? &c.tm_wf_half : nullptr, c.win_fwd ? &c.tm_in_win : nullptr, a, st);
        if (rc) return rc;
    }
    return AG_OK;
}

// Forward pass (policy.py:26-42 `forward`).  d_planes uint8 [n][in_planes][board][board].
// policy head: d_out fp32 [n][board*board] probabilities; value head: d_out fp32 [n] in [-1, 1].
int ag_net_forward(ag_net* net, const uint8_t* d_planes, int n, float* d_out, void* stream) {
    if (n < 0 || n > net->max_batch) { set_error("ag_net_forward: batch %d exceeds max_batch %d", n, net->max_batch); return AG_EINVAL; }
    if (n == 0) return AG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int rc = net_conv_stack(net, d_planes, n, st);
    if (rc) return rc;
    ConvLayer& cl = net->L[net->n_conv - 1];
    if (net->head_kind == 0) {
        launch_k(policy_head_kernel, dim3(n), dim3(kHeadThreads), 0, st, cl.act, cl.ldout, cl.lout, net->hw, net->hb, net->use_pos_bias ? net->hpos : nullptr,
                                                      d_out, nullptr, 0.f, nullptr, nullptr, nullptr, nullptr, 0);
        AG_LAUNCH_CHECK("policy_head_kernel");
    } else {
        launch_k(value_head_kernel, dim3(n), dim3(256), 0, st, cl.act, cl.ldout, cl.lout, net->hw, net->hb, net->d1w, net->d1b, net->d2w, net->d2b,
                                             net->hidden, d_out);
        AG_LAUNCH_CHECK("value_head_kernel");
    }
    return AG_OK;
}

// Forward + fused softmax / categorical cross-entropy / accuracy head and its backward
// (spt.py:200 compile(loss='categorical_crossentropy', metrics=['accuracy'])).  d_labels int32 [n] =
// flat move index x*board+y.  d_loss_acc fp32[2] is ACCUMULATED into: sum of losses, number correct.
// Leaves the gradient w.r.t. the last conv layer's pre-activation in place for ag_net_backward_layer.
// d_probs may be NULL.  Gradients are means over this call's n.
int ag_net_forward_loss(ag_net* net, const uint8_t* d_planes, const int32_t* d_labels, int n, float* d_probs, float* d_loss_acc, void* stream) {
    if (net->head_kind != 0) { set_error("ag_net_forward_loss: policy head only"); return AG_ESTATE; }
    if (n < 1 || n > net->max_batch) { set_error("ag_net_forward_loss: batch %d not in 1..%d", n, net->max_batch); return AG_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    int rc = net_conv_stack(net, d_planes, n, st);
    if (rc) return rc;
    rc = net_build_wgrad_maps(net, n);
    if (rc) return rc;
    net->dgrad_chain_done = false;
    const int last = net->n_conv - 1;
    ConvLayer& cl = net->L[last];
    bf16* G = (last == 0 && net->x0_grad) ? net->x0_grad : net->gbuf[last];
    if (last == 0 && net->x0_grad) { set_error("single-layer nets with a wide first filter are not trainable"); return AG_ESTATE; }
    launch_k(policy_head_kernel, dim3(n), dim3(kHeadThreads), 0, st, cl.act, cl.ldout, cl.lout, net->hw, net->hb, net->use_pos_bias ? net->hpos : nullptr,
                                                  d_probs, d_labels, 1.f / n, G, net->head_partial, d_loss_acc, nullptr, net->loss_kind);
    AG_LAUNCH_CHECK("policy_head_kernel");
    launch_k(head_reduce_kernel, dim3((625 + 7) / 8), dim3(256), 0, st, net->head_partial, n, cl.ldout, net->board * net->board, net->g_hw, net->g_hb,
                                                        net->use_pos_bias ? net->g_hpos : nullptr);
    AG_LAUNCH_CHECK("head_reduce_kernel");
    return AG_OK;
}

// weight (+ bias) gradient of one layer: its own kernel + the split reduction kernel
static int net_wgrad_one(ag_net* net, int layer, int n, cudaStream_t st) {
    ConvLayer& c = net->L[layer];
    const bf16* G = (layer == 0 && net->x0_grad) ? net->x0_grad : net->gbuf[layer];
    const size_t rows = (size_t)n * c.lin.img_stride;
    WgradArgs w;
    memset(&w, 0, sizeof(w));
    w.m_total = (int)rows; w.taps = c.taps; memcpy(w.off, c.off_fwd, sizeof(int) * c.taps);
    w.ldout = c.ldout; w.kin = c.kin; w.splits = c.wgrad_splits; w.scratch = net->wgrad_scratch;
    w.kb_img = wgrad_kb_img(c.lin, c.tm_wg.krows); w.img_stride = c.lin.img_stride;
    const bool fused_bias = c.kin + 16 <= 256;  // the bias gradient rides in the weight-gradient GEMM as an extra column of ones
    w.bias_scratch = fused_bias ? net->wgrad_scratch + net->wgrad_scratch_floats : nullptr;
    int rc = wgrad_launch(c.tm_wg, w, c.gw, fused_bias ? c.gb : nullptr, st);
    if (rc) return rc;
    if (!fused_bias) {
        launch_k(colsum_stage1_kernel, dim3(net->colsum_blocks), dim3(256), 0, st, G, rows, c.ldout, net->colsum_part);
        AG_LAUNCH_CHECK("colsum_stage1_kernel");
        launch_k(colsum_stage2_kernel, dim3((c.ldout + 7) / 8), dim3(256), 0, st, net->colsum_part, net->colsum_blocks, c.ldout, c.ldout, c.gb);
        AG_LAUNCH_CHECK("colsum_stage2_kernel");
    }
    return AG_OK;
}

static bool same_wgrad_geometry(const ConvLayer& a, const ConvLayer& b) {
    return a.taps == b.taps && a.ldout == b.ldout && a.kin == b.kin && a.wgrad_splits == b.wgrad_splits && a.lin.img_stride == b.lin.img_stride &&
           a.tm_wg.win_rows == b.tm_wg.win_rows && a.tm_wg.krows == b.tm_wg.krows && memcmp(a.off_fwd, b.off_fwd, sizeof(int) * a.taps) == 0;
}

// weight gradients of layers hi .. lo: runs of layers with one geometry go out as chained launches (wgrad.cu)
static int net_wgrad_range(ag_net* net, int hi, int lo, int n, cudaStream_t st) {
    // The first layer's weight gradient needs only the data-gradient chain's result, and its own kernels do not join the
    // chained launch: queued FIRST, its gradient can be exchanged while the chain runs, and what is left exposed at the
    // end of the backward pass is the exchange of the chain's last layer instead of two buckets.  An event marks the point.
    net->first_layer_early = false;
    if (lo == 0 && hi > 0 && !net->L[0].wgrad_chain && net->L[1].wgrad_chain) {
        int rc = net_wgrad_one(net, 0, n, st);
        if (rc) return rc;
        if (!net->ev_first) AG_CUDA(cudaEventCreateWithFlags(&net->ev_first, cudaEventDisableTiming));
        AG_CUDA(cudaEventRecord(net->ev_first, st));
        net->first_layer_early = true;
        lo = 1;
    }
    int l = hi;
    while (l >= lo) {
        ConvLayer& first = net->L[l];
        if (!first.wgrad_chain) {
            int rc = net_wgrad_one(net, l, n, st);
            if (rc) return rc;
            l--;
            continue;
        }
        WgradChainEntry e[kWgradChainMaxLayers];
        int cnt = 0;
        while (l >= lo && cnt < kWgradChainMaxLayers && net->L[l].wgrad_chain && same_wgrad_geometry(net->L[l], first)) {
            ConvLayer& c = net->L[l];
            float* scratch = (cnt & 1) ? net->wgrad_scratch2 : net->wgrad_scratch;
            e[cnt].maps = &c.tm_wg;
            e[cnt].scratch = scratch;
            e[cnt].bias_scratch = scratch + net->wgrad_scratch_floats;
            e[cnt].dw = c.gw; e[cnt].db = c.gb;
            e[cnt].counters = net->wgrad_counters + (size_t)l * kWgradCounterStride;
            e[cnt].target = ++c.wgrad_chain_count * (unsigned int)c.wgrad_splits;
            e[cnt].done = e[cnt].counters + (kWgradCounterStride - 1);   // the layer's last counter slot (tiles < stride)
            cnt++;
            l--;
        }
        WgradArgs w;
        memset(&w, 0, sizeof(w));
        w.m_total = n * first.lin.img_stride; w.taps = first.taps; memcpy(w.off, first.off_fwd, sizeof(int) * first.taps);
        w.ldout = first.ldout; w.kin = first.kin; w.splits = first.wgrad_splits;
        w.kb_img = wgrad_kb_img(first.lin, first.tm_wg.krows); w.img_stride = first.lin.img_stride;
        const unsigned int ctas = (unsigned int)wgrad_chain_ctas(w);
        for (int k = l + cnt; k > l; k--) net->L[k].wgrad_done_target = net->L[k].wgrad_chain_count * ctas;
        int rc = wgrad_chain_launch(w, e, cnt, st);
        if (rc == AG_ESTATE) {   // refused (grid not co-resident): these and all later layers use their own kernel + reduce
            net->wgrad_chain_refused = true;
            net->chain_fallbacks++;
            for (auto& c : net->L) c.wgrad_chain = false;
            for (int k = l + cnt; k > l; k--) {
                rc = net_wgrad_one(net, k, n, st);
                if (rc) return rc;
            }
            continue;
        }
        if (rc) return rc;
    }
    return AG_OK;
}

// Backward of conv layers hi .. lo (hi >= lo; call with descending, non-overlapping ranges after ag_net_forward_loss,
// starting at n_conv-1): weight + bias gradients into the flat gradient buffer and the data gradients into the layers
// below, each gated by the ReLU of the layer it lands in.
int ag_net_backward_range(ag_net* net, int hi, int lo, int n, void* stream) {
    if (lo < 0 || hi >= net->n_conv || lo > hi || n < 1 || n > net->max_batch || net->wgrad_n != n) {
        set_error("ag_net_backward_range: bad layers/batch or no forward_loss before it");
        return AG_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (net->chain_bwd) {
        if (!net->dgrad_chain_done) {
            // every data gradient of this backward pass in one launch (G of every layer has its own buffer); the weight
            // gradients follow range by range as the caller asks for them
            rc = net_build_chains(net, n);
            if (rc) return rc;
            ++net->chain_epoch_bwd;
            for (const ChainParams& P : net->chain_bwd_p) {
                rc = conv_chain_launch(P, net->chain_flags_bwd, net->chain_epoch_bwd, net->chain_win_rows, net->chain_nmax_bwd, st);
                if (rc == AG_ESTATE) break;
                if (rc) return rc;
            }
            if (rc == AG_ESTATE) {   // refused: the data gradients go layer by layer, this range now and the others as they come
                if (hi != net->n_conv - 1) { set_error("data-gradient chain refused after part of the backward pass was queued"); return AG_ESTATE; }
                net->chain_bwd = false;
                net->chain_fallbacks++;
            } else {
                net->dgrad_chain_done = true;
            }
        }
    }
    if (!net->chain_bwd) {
        for (int l = hi; l >= lo && l > 0; l--) {  // data gradient of layer l into the layer below (needs only G of layer l)
            ConvLayer& c = net->L[l];
            ConvArgs a;
            bwd_conv_args(net, l, n, &a);
            rc = conv_gemm_launch(c.tm_g, c.tm_wb, c.pair_bwd ? &c.tm_wb_half : nullptr, c.win_bwd ? &c.tm_g_win : nullptr, a, st);
            if (rc) return rc;
        }
    }
    return net_wgrad_range(net, hi, lo, n, st);
}

// Backward of one conv layer (call for layer = n_conv-1 down to 0 after ag_net_forward_loss).
int ag_net_backward_layer(ag_net* net, int layer, int n, void* stream) { return ag_net_backward_range(net, layer, layer, n, stream); }

// Flat fp32 gradient buffer (device): conv layers in order (weights [tap][ldout][kin] then bias [ldout]),
// then the head (w[256], b[1], pos[368]).  h_offsets receives n_conv+2 element offsets (layer starts,
// head start, end).  This is what data-parallel training all-reduces.
int ag_net_grads(ag_net* net, float** d_grads, long long* count, long long* h_offsets) {
    *d_grads = net->grads;
    *count = (long long)net->grad_count;
    if (h_offsets) for (size_t i = 0; i < net->grad_offsets.size(); i++) h_offsets[i] = net->grad_offsets[i];
    return AG_OK;
}

// Fused SGD(+momentum) update over every parameter (spt.py:199 SGD(lr, decay); the host applies Keras'
// lr/(1+decay*iterations) schedule and passes the step's lr).  grad_scale multiplies the gradients
// first (1/world_size after a sum all-reduce).  Also refreshes the bf16 operand copies.
int ag_net_sgd(ag_net* net, float lr, float momentum, float grad_scale, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    SgdFlats flats;
    memset(&flats, 0, sizeof(flats));
    flats.f[0] = SgdFlat{net->hw, net->vhw, net->g_hw, 256};
    flats.f[1] = SgdFlat{net->hb, net->vhb, net->g_hb, 1};
    flats.n = 2;
    if (net->use_pos_bias) flats.f[flats.n++] = SgdFlat{net->hpos, net->vhpos, net->g_hpos, net->board * net->board};
    const int total = net->sgd_tiles + flats.n;
    launch_k(sgd_all_kernel, dim3(total < sm_count() * 8 ? total : sm_count() * 8), dim3(256), 0, st, (const SgdLayer*)net->sgd_table, net->n_conv,
             net->sgd_tiles, flats, lr, momentum, grad_scale);
    AG_LAUNCH_CHECK("sgd_all_kernel");
    net->iterations++;
    return AG_OK;
}

// Where a chained layer's gradient becomes final while the launch is still running: the device address of a counter and the
// value it reaches when dw / db of `layer` from the latest ag_net_backward_range are complete.  AG_ESTATE when the layer's
// weight gradient did not run chained (its gradient is final when the backward call's kernels are: wait for the stream).
int ag_net_wgrad_done(ag_net* net, int layer, unsigned int** d_counter, unsigned int* target) {
    if (layer < 0 || layer >= net->n_conv) { set_error("ag_net_wgrad_done: bad layer %d", layer); return AG_EINVAL; }
    ConvLayer& c = net->L[layer];
    if (!c.wgrad_chain || c.wgrad_chain_count == 0) { set_error("ag_net_wgrad_done: layer %d does not run chained", layer); return AG_ESTATE; }
    *d_counter = net->wgrad_counters + (size_t)layer * kWgradCounterStride + (kWgradCounterStride - 1);
    *target = c.wgrad_done_target;
    return AG_OK;
}

// Makes `stream` wait for the first layer's weight gradient of the latest ag_net_backward_range when that call queued it ahead
// of the chained launch (AG_OK), else AG_ESTATE: the gradient is then final where the backward call ends in its stream.
int ag_net_wait_first_layer(ag_net* net, void* stream) {
    if (!net->first_layer_early || !net->ev_first) { set_error("ag_net_wait_first_layer: the first layer was not queued early"); return AG_ESTATE; }
    AG_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, net->ev_first, 0));
    return AG_OK;
}

// Makes `stream` wait until the 32-bit word at d_addr has reached `value` (cuStreamWaitValue32, >= in wrap-around
// arithmetic): everything queued on the stream afterwards starts only then.
int ag_stream_wait_value(void* stream, const unsigned int* d_addr, unsigned int value) {
    typedef CUresult (*wait_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
    static wait_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) { set_error("cuStreamWaitValue32 is not available"); return AG_ECUDA; }
        fn = (wait_fn)p;
    }
    CUresult r = fn((CUstream)stream, (CUdeviceptr)(uintptr_t)d_addr, value, CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) { set_error("cuStreamWaitValue32 failed (%d)", (int)r); return AG_ECUDA; }
    return AG_OK;
}

// SMs the weight-gradient launches may fill (0 = all, the default).  Data-parallel training sets this a little below the SM
// count: the chained weight gradients need all their CTAs resident at once, and the all-reduce of the previous gradient
// bucket is running on a few SMs at the same time.
int ag_net_set_sm_budget(ag_net* net, int sms) {
    if (sms < 0) { set_error("ag_net_set_sm_budget: negative"); return AG_EINVAL; }
    net->sm_budget = sms;
    net->wgrad_n = -1;  // re-plan the split factors at the next forward_loss
    return AG_OK;
}

// Loss the head kernel differentiates: 0 = categorical cross-entropy (supervised_policy_trainer.py:200), 1 = Keras'
// binary_crossentropy applied to the 361 softmax outputs (reinforcement_policy_trainer.py:109).  d_loss_acc[0] accumulates
// that loss; [1] stays the argmax-match count.
int ag_net_set_loss(ag_net* net, int kind) {
    if (kind != 0 && kind != 1) { set_error("ag_net_set_loss: kind %d (0 categorical, 1 binary cross-entropy)", kind); return AG_EINVAL; }
    net->loss_kind = kind;
    return AG_OK;
}

// One single-GPU SL training step: forward, loss, full backward, SGD update.
int ag_net_train_step(ag_net* net, const uint8_t* d_planes, const int32_t* d_labels, int n, float lr, float momentum,
                      float* d_loss_acc, void* stream) {
    int rc = ag_net_forward_loss(net, d_planes, d_labels, n, nullptr, d_loss_acc, stream);
    if (!rc) rc = ag_net_backward_range(net, net->n_conv - 1, 0, n, stream);
    if (!rc) rc = ag_net_sgd(net, lr, momentum, 1.f, stream);
    return rc;
}

// Measurement hook: re-runs ONE kernel of the step on the buffers the last forward_loss/backward left in
// place: mode 0 = forward conv of `layer`, 1 = its data gradient, 2 = its weight gradient (+ reduce), 3 = the chained
// forward stack (all layers, one launch), 4 = the chained data gradients.
int ag_net_run_layer(ag_net* net, int layer, int mode, int n, void* stream) {
    if (layer < 0 || layer >= net->n_conv || n < 1 || n > net->max_batch) { set_error("ag_net_run_layer: bad argument"); return AG_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    ConvLayer& c = net->L[layer];
    if (mode == 0) {
        ConvArgs a;
        fwd_conv_args(net, layer, n, &a);
        return conv_gemm_launch(c.tm_in, c.tm_wf, c.pair_fwd ? &c.tm_wf_half : nullptr, c.win_fwd ? &c.tm_in_win : nullptr, a, st);
    }
    if (mode == 3 || mode == 4) {  // the whole forward stack / data-gradient chain as the step launches it (`layer` ignored)
        if (!(mode == 3 ? net->chain_fwd : net->chain_bwd) || net->chain_n != n) { set_error("ag_net_run_layer: no layer chain built for this batch"); return AG_ESTATE; }
        unsigned int epoch = mode == 3 ? ++net->chain_epoch_fwd : ++net->chain_epoch_bwd;
        for (const ChainParams& P : (mode == 3 ? net->chain_fwd_p : net->chain_bwd_p)) {
            int rc = conv_chain_launch(P, mode == 3 ? net->chain_flags_fwd : net->chain_flags_bwd, epoch, net->chain_win_rows,
                                       mode == 3 ? net->chain_nmax_fwd : net->chain_nmax_bwd, st);
            if (rc) return rc;
        }
        return AG_OK;
    }
    if (net->wgrad_n != n) { set_error("ag_net_run_layer: run ag_net_forward_loss with this batch first"); return AG_ESTATE; }
    if (mode == 5) {  // the weight gradients of every hidden layer as the step launches them: one chained launch (`layer` ignored)
        if (net->n_conv < 3) { set_error("ag_net_run_layer: no hidden layers"); return AG_EINVAL; }
        for (int l = 1; l + 1 < net->n_conv; l++)
            if (!net->L[l].wgrad_chain) { set_error("ag_net_run_layer: layer %d does not run chained", l); return AG_ESTATE; }
        return net_wgrad_range(net, net->n_conv - 2, 1, n, st);
    }
    if (mode == 2) {
        WgradArgs w;
        memset(&w, 0, sizeof(w));
        w.m_total = n * c.lin.img_stride; w.taps = c.taps; memcpy(w.off, c.off_fwd, sizeof(int) * c.taps);
        w.ldout = c.ldout; w.kin = c.kin; w.splits = c.wgrad_splits; w.scratch = net->wgrad_scratch;
        w.bias_scratch = c.kin + 16 <= 256 ? net->wgrad_scratch + net->wgrad_scratch_floats : nullptr;
        w.kb_img = wgrad_kb_img(c.lin, c.tm_wg.krows); w.img_stride = c.lin.img_stride;
        return wgrad_launch(c.tm_wg, w, c.gw, w.bias_scratch ? c.gb : nullptr, st);
    }
    if (mode == 1 && layer > 0) {
        ConvArgs a;
        bwd_conv_args(net, layer, n, &a);
        return conv_gemm_launch(c.tm_g, c.tm_wb, c.pair_bwd ? &c.tm_wb_half : nullptr, c.win_bwd ? &c.tm_g_win : nullptr, a, st);
    }
    set_error("ag_net_run_layer: mode %d not available for layer %d", mode, layer);
    return AG_EINVAL;
}

// Test hook: `ctas` CTAs that each take an SM's worth of shared memory and sit there for `microseconds` (globaltimer), on
// `stream`.  Lets a test put the chained launches next to a kernel that holds SMs (what another net, NCCL or a torch op on
// a second stream does in production).  Bounded: at most 5 s.
int ag_debug_hold_sms(int ctas, long long microseconds, void* stream) {
    if (ctas < 1 || ctas > 1024 || microseconds < 0 || microseconds > 5000000) { set_error("ag_debug_hold_sms: bad argument"); return AG_EINVAL; }
    static PerDevice<bool> configured;
    const int smem = 160 * 1024;   // more than half an SM: one holder per SM
    if (!configured.cur()) {
        AG_CUDA(cudaFuncSetAttribute(hold_sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured.cur() = true;
    }
    hold_sm_kernel<<<ctas, 32, smem, (cudaStream_t)stream>>>(microseconds * 1000);
    AG_LAUNCH_CHECK("hold_sm_kernel");
    return AG_OK;
}

// Test/debug readback (synchronous): output activations of conv layer `layer` (kind 0) or the
// pre-activation gradient currently held for it (kind 1) as fp32 [n][cout][board][board] on the host.
int ag_net_debug_read(ag_net* net, int layer, int kind, int n, float* h_out) {
    if (layer < 0 || layer >= net->n_conv || n < 1 || n > net->max_batch) { set_error("ag_net_debug_read: bad argument"); return AG_EINVAL; }
    ConvLayer& c = net->L[layer];
    const bf16* src; PixelLayout lay;
    if (kind == 0) { src = c.act; lay = c.lout; }
    else if (layer == 0 && net->x0_grad) { src = net->x0_grad; lay = net->lay_in; }
    else { src = net->gbuf[layer]; lay = c.lout; }
    size_t count = (size_t)n * c.cout * net->board * net->board;
    float* d = nullptr;
    AG_CUDA(cudaMalloc(&d, count * 4));
    unpack_rows_kernel<<<grid_for(count, 256), 256>>>(src, n, c.cout, net->board, lay, c.ldout, d);
    count_launch();
    cudaError_t e = cudaMemcpy(h_out, d, count * 4, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return cuda_fail(e, "ag_net_debug_read");
    return AG_OK;
}

}  // extern "C"
