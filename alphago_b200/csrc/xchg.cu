// Gradient exchange over NVLink peer memory: a two-shot sum all-reduce written for the data-parallel SL step
// ("one exchange step"; the reference trains on one device and has no counterpart).
//
// Every rank of the node maps every other rank's gradient buffer (CUDA IPC).  One kernel per bucket, a few CTAs:
//   0. tell every peer "my part of this bucket is final" and wait until all of them have said so
//   1. reduce-scatter: this rank owns slice r of the bucket and sums it over all ranks' buffers with plain 16-byte peer
//      loads, always in rank order 0..world-1, into its own buffer
//   2. tell every peer "my slice is reduced", wait for theirs, and copy the other slices from their owners
//   3. count itself off in every rank's signal block ("I have stopped reading"); a rank waits for that count -- a stream
//      memory operation, ag_xchg_wait_readers -- before its next backward pass writes its buffer again
// Flags are 32-bit sequence numbers in each rank's own signal block, written by the peers with system-scope release
// stores and polled locally with acquire loads; they only ever grow, so a rank that runs ahead cannot confuse a slower one.
// Because the owner alone reduces a slice and everybody copies its result, all ranks end with bit-identical sums
// (a ring all-reduce gives each rank its own summation order).  Measured against NCCL on this box (8 x B200 without NVLS:
// ring, LL protocol, 118 us for the 15.5 MB of one step): see DESIGN.md section 5.
//
// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung
// the GPU; this kernel reads peers with ordinary ld.global only.
#include <cuda.h>
#include <string.h>

#include "common.cuh"

extern "C" int ag_stream_wait_value(void* stream, const unsigned int* d_addr, unsigned int value);

namespace ag {

constexpr int kXchgMaxWorld = 16;
constexpr int kXchgThreads = 256;   // x 77 registers: a CTA fits next to a chained CTA of the network on the same SM

struct XchgPeers {
    float* buf[kXchgMaxWorld];            // every rank's gradient buffer (own one included), same layout
    unsigned int* sig[kXchgMaxWorld];     // every rank's signal block: [3][kXchgMaxWorld] words + CTA counters
};
// signal block layout (words): ready[w] at 0, reduced[w] at 16, finished[w] at 32, CTA arrival counters at 48, 49
constexpr int kSigReady = 0, kSigReduced = 16, kSigFinished = 32, kSigCtas1 = 48, kSigCtas2 = 49, kSigWords = 64;

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// wait until all `world` words of my own block starting at `base` have reached seq (wrap-around compare); one lane per word
__device__ __forceinline__ void wait_all(const unsigned int* mine, int base, int world, unsigned int seq) {
    if (threadIdx.x < (unsigned)world) {
        long long t0 = 0;
        while ((int)(ld_acquire_sys(mine + base + threadIdx.x) - seq) < 0) {
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > 20000000000LL) {   // ~10 s: a peer died; do not hang the device for ever
                printf("alphago_b200: gradient exchange timed out waiting for rank %d (flag %d, seq %u)\n", (int)threadIdx.x, base, seq);
                __trap();
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kXchgThreads)
xchg_allreduce_kernel(const XchgPeers P, int rank, int world, size_t start, size_t count, unsigned int seq) {
    unsigned int* mine = P.sig[rank];
    // ---- 0. my bucket is final (stream order / the caller's gate says so): tell everybody, wait for everybody
    if (blockIdx.x == 0 && threadIdx.x < (unsigned)world) {
        __threadfence_system();
        st_release_sys(P.sig[threadIdx.x] + kSigReady + rank, seq);
    }
    wait_all(mine, kSigReady, world, seq);

    // ---- 1. reduce my slice (count is a multiple of 4 floats, slices are too)
    const size_t n4 = count / 4;
    const size_t per = (n4 + world - 1) / world;
    const size_t lo = per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    // kU independent 16-byte groups per thread and pass, every load issued before the first add: the loop is bound by NVLink
    // latency (2-3 us a round trip), so what counts is how many loads are in flight (one group per pass took 120 us a bucket)
    constexpr int kU = 2, kR = 8;   // 16 loads in flight per thread: two groups x eight ranks
    for (size_t i0 = lo + tid; i0 < hi; i0 += nthr * kU) {
        float4 a[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
        for (int r0 = 0; r0 < world; r0 += kR) {
            float4 v[kR][kU];
#pragma unroll
            for (int rr = 0; rr < kR; rr++) {
                const float4* src = reinterpret_cast<const float4*>(P.buf[r0 + rr < world ? r0 + rr : rank] + start);
#pragma unroll
                for (int u = 0; u < kU; u++) {
                    const size_t i = i0 + (size_t)u * nthr;
                    v[rr][u] = (r0 + rr < world && i < hi) ? __ldcv(src + i) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
#pragma unroll
            for (int rr = 0; rr < kR; rr++)       // rank order 0, 1, 2, ...: the same sum on whichever rank owns the slice
#pragma unroll
                for (int u = 0; u < kU; u++) { a[u].x += v[rr][u].x; a[u].y += v[rr][u].y; a[u].z += v[rr][u].z; a[u].w += v[rr][u].w; }
        }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const size_t i = i0 + (size_t)u * nthr;
            if (i < hi) reinterpret_cast<float4*>(P.buf[rank] + start)[i] = a[u];
        }
    }
    // all my CTAs done -> the last one tells the peers
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned int arrived = atomicAdd(mine + kSigCtas1, 1u) + 1u;
        if (arrived == gridDim.x * seq) {
            for (int r = 0; r < world; r++) st_release_sys(P.sig[r] + kSigReduced + rank, seq);
        }
    }
    // ---- 2. fetch the other slices from their owners
    wait_all(mine, kSigReduced, world, seq);
    {
        // one flat index over (owner, element) so that loads from ALL owners are in flight together (owner by owner, each pass
        // paid an NVLink round trip: 7 of them on 8 GPUs); thread t starts with owner rank + 1 + t % (world - 1)
        const size_t others = (size_t)(world - 1) * per;
        float4* dst = reinterpret_cast<float4*>(P.buf[rank] + start);
        constexpr int kC = 8;
        for (size_t j0 = tid; j0 < others; j0 += nthr * kC) {
            float4 v[kC];
            size_t at[kC];
#pragma unroll
            for (int u = 0; u < kC; u++) {
                const size_t j = j0 + (size_t)u * nthr;
                at[u] = (size_t)-1;
                if (j < others) {
                    const int r = (rank + 1 + (int)(j % (size_t)(world - 1))) % world;   // interleave the owners
                    const size_t i = per * r + j / (size_t)(world - 1);
                    if (i < n4 && i < per * (r + 1)) { at[u] = i; v[u] = __ldcv(reinterpret_cast<const float4*>(P.buf[r] + start) + i); }
                }
            }
#pragma unroll
            for (int u = 0; u < kC; u++) if (at[u] != (size_t)-1) dst[at[u]] = v[u];
        }
    }
    // ---- 3. I have stopped reading: count myself off in every rank's block.  Nobody waits here: the owner of a buffer waits for
    // the count (ag_xchg_wait_readers, a stream memory operation) before its next backward pass writes the buffer again
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned int arrived = atomicAdd(mine + kSigCtas2, 1u) + 1u;
        if (arrived == gridDim.x * seq) {
            for (int r = 0; r < world; r++) asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(P.sig[r] + kSigFinished) : "memory");
        }
    }
}

}  // namespace ag

struct ag_xchg {
    int rank, world, ctas;
    float* local;
    size_t count;
    unsigned int* sig;           // my signal block (device)
    ag::XchgPeers peers;
    void* opened[2 * ag::kXchgMaxWorld];
    int n_opened;
    unsigned int seq;
    long long buf_offset;        // of `local` inside its allocation (IPC handles name allocations)
};

extern "C" {

// Creates this rank's end of the exchange over d_buf[count] (the buffer ag_net_grads() returns).  h_handle receives 144 bytes
// to be all-gathered by the caller: the IPC handles of the buffer's allocation and of the signal block, and the buffer's
// offset inside its allocation.
ag_xchg* ag_xchg_create(int rank, int world, float* d_buf, long long count, int ctas, unsigned char* h_handle) {
    using namespace ag;
    if (world < 2 || world > kXchgMaxWorld || rank < 0 || rank >= world || !d_buf || count < 4 || ctas < 1 || ctas > 148) {
        set_error("ag_xchg_create: bad argument (world 2..%d)", kXchgMaxWorld);
        return nullptr;
    }
    ag_xchg* x = new ag_xchg();
    memset(x, 0, sizeof(*x));
    x->rank = rank; x->world = world; x->ctas = ctas; x->local = d_buf; x->count = (size_t)count;
    if (cudaMalloc(&x->sig, kSigWords * sizeof(unsigned int)) != cudaSuccess || cudaMemset(x->sig, 0, kSigWords * sizeof(unsigned int)) != cudaSuccess) {
        set_error("ag_xchg_create: cannot allocate the signal block");
        delete x;
        return nullptr;
    }
    CUdeviceptr base = 0;
    size_t size = 0;
    typedef CUresult (*range_fn)(CUdeviceptr*, size_t*, CUdeviceptr);
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fp, cudaEnableDefault, &q) != cudaSuccess || !fp ||
        ((range_fn)fp)(&base, &size, (CUdeviceptr)(uintptr_t)d_buf) != CUDA_SUCCESS) {
        set_error("ag_xchg_create: cuMemGetAddressRange failed");
        cudaFree(x->sig); delete x;
        return nullptr;
    }
    x->buf_offset = (long long)((uintptr_t)d_buf - (uintptr_t)base);
    cudaIpcMemHandle_t hb, hs;
    if (cudaIpcGetMemHandle(&hb, (void*)(uintptr_t)base) != cudaSuccess || cudaIpcGetMemHandle(&hs, x->sig) != cudaSuccess) {
        set_error("ag_xchg_create: cudaIpcGetMemHandle failed (%s)", cudaGetErrorString(cudaGetLastError()));
        cudaFree(x->sig); delete x;
        return nullptr;
    }
    memcpy(h_handle, &hb, 64);
    memcpy(h_handle + 64, &hs, 64);
    memcpy(h_handle + 128, &x->buf_offset, 8);
    memset(h_handle + 136, 0, 8);
    return x;
}

// h_all: the 144-byte records of all ranks in rank order.  Maps the peers' buffers and signal blocks.
int ag_xchg_connect(ag_xchg* x, const unsigned char* h_all) {
    using namespace ag;
    for (int r = 0; r < x->world; r++) {
        if (r == x->rank) { x->peers.buf[r] = x->local; x->peers.sig[r] = x->sig; continue; }
        cudaIpcMemHandle_t hb, hs;
        long long off;
        memcpy(&hb, h_all + (size_t)r * 144, 64);
        memcpy(&hs, h_all + (size_t)r * 144 + 64, 64);
        memcpy(&off, h_all + (size_t)r * 144 + 128, 8);
        void *pb = nullptr, *ps = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&pb, hb, cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess) { x->opened[x->n_opened++] = pb; e = cudaIpcOpenMemHandle(&ps, hs, cudaIpcMemLazyEnablePeerAccess); }
        if (e != cudaSuccess) { (void)cudaGetLastError(); set_error("ag_xchg_connect: cannot map rank %d's memory (%s)", r, cudaGetErrorString(e)); return AG_ECUDA; }
        x->opened[x->n_opened++] = ps;
        x->peers.buf[r] = reinterpret_cast<float*>(static_cast<unsigned char*>(pb) + off);
        x->peers.sig[r] = static_cast<unsigned int*>(ps);
    }
    return AG_OK;
}

// In-place sum all-reduce of buf[start : start + count) on every rank (same arguments, same call order on all ranks; start and
// count multiples of 4 floats).  Everything queued on `stream` before the call must make this rank's part of the range final.
int ag_xchg_allreduce(ag_xchg* x, long long start, long long count, void* stream) {
    using namespace ag;
    if (start < 0 || count < 0 || (start & 3) || (count & 3) || (size_t)(start + count) > x->count) {
        set_error("ag_xchg_allreduce: range [%lld, +%lld) must be 4-float aligned and inside the buffer of %zu", start, count, x->count);
        return AG_EINVAL;
    }
    if (count == 0) return AG_OK;
    x->seq++;
    xchg_allreduce_kernel<<<x->ctas, kXchgThreads, 0, (cudaStream_t)stream>>>(x->peers, x->rank, x->world, (size_t)start, (size_t)count, x->seq);
    AG_LAUNCH_CHECK("xchg_allreduce_kernel");
    return AG_OK;
}

// Makes `stream` wait until every rank has finished reading this rank's buffer in all exchanges issued so far: call it on
// the stream that will overwrite the buffer (the next backward pass) before it does.  Costs no SM time.
int ag_xchg_wait_readers(ag_xchg* x, void* stream) {
    if (x->seq == 0) return AG_OK;
    return ag_stream_wait_value(stream, x->sig + ag::kSigFinished, (unsigned int)x->world * x->seq);
}

void ag_xchg_destroy(ag_xchg* x) {
    if (!x) return;
    cudaDeviceSynchronize();
    for (int i = 0; i < x->n_opened; i++) cudaIpcCloseMemHandle(x->opened[i]);
    cudaFree(x->sig);
    delete x;
}

}  // extern "C"
