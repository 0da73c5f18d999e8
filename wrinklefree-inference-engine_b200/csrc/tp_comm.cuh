// Tensor-parallel exchange over NVLink peer memory (north-star item 6).
//
// One process per GPU.  Every rank owns one "comm" allocation (cudaMalloc, exported with cudaIpcGetMemHandle and
// opened by the peers) that holds the buffers other ranks write INTO:
//   attn_full  bf16 [MB][H*D]      all-gather target of the attention outputs  (each rank stores its heads into all)
//   h_full     bf16 [MB][ffn]      all-gather target of relu^2(gate)*up
//   acc_o      f32  [MB][hidden]   all-reduce target of the o-proj partial sums    (each rank REDs into all)
//   acc_down   f32  [MB][hidden]   all-reduce target of the down-proj partial sums
//   cand       {f32,i32} [MB][8]   per-rank argmax candidates of the vocab-split LM head
//   flags      i32  [n_flags][8]   flags[id][r] = last step in which rank r finished exchange `id`
// The collective is fused into the producing kernel: attention / linear CTAs store or RED straight into every
// peer's buffer (st/red.global over NVLink), and the LAST CTA of the grid publishes the rank's flag to all peers
// (membar.sys + st.release.sys).  The consuming kernel's prologue spins on its LOCAL flags (ld.acquire.sys).
// All values exchanged by RED are integers < 2^24 held in fp32, so the all-reduce is exact and order-independent:
// TP = N is bit-identical to TP = 1 (every rank quantises the gathered full vectors exactly like TP = 1 does).
// Flag values are the 1-based decode step (a device-side counter the graph-captured step increments itself), so a
// captured CUDA graph can be replayed without patching kernel parameters.
#pragma once
#include "common.cuh"

namespace wf {

constexpr int TP_MAX = 8;

struct TpComm {
    int rank, size;
    unsigned char* peer[TP_MAX];   // base of every rank's comm buffer in THIS process' address space (peer[rank] = local)
    size_t off_flags;              // byte offsets inside a comm buffer
    const int* step;               // local device counter: number of completed decode steps
    int* done_cnt;                 // local, zero-initialised: one grid-arrival counter per flag id
    int* err;                      // local device word: set non-zero when a bounded spin gave up (checked by the host at every sync)
};

// Every device-side spin in this library is bounded: a peer that died, or ranks whose step counters disagree, must become an
// error code on the host, not a hung GPU.  The budget is generous (seconds) against waits that are microseconds when healthy.
constexpr unsigned long long WF_SPIN_BUDGET_NS = 4000000000ull;
__device__ __forceinline__ unsigned long long spin_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ int* tp_flag_ptr(const TpComm& c, int dst_rank, int id, int src_rank) {
    return reinterpret_cast<int*>(c.peer[dst_rank] + c.off_flags) + id * TP_MAX + src_rank;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_relaxed_gpu(const int* p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Called by EVERY thread of EVERY CTA of the producing grid after its last peer store / RED.  The CTA that arrives
// last publishes flag `id` = step + 1 on all ranks.  `nblocks` = total CTAs of the grid.
__device__ __forceinline__ void tp_signal_grid_done(const TpComm& c, int id, int nblocks) {
    if (c.size <= 1) return;
    __syncthreads();                 // every thread's peer stores / REDs happen-before warp 0's fences (CTA barrier) ...
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int last = 0;
        if (lane == 0) {
            __threadfence_system();  // ... which make them visible system-wide before the CTA's arrival (ONE membar.sys per CTA:
                                     // a fence per thread serialises tens of thousands of system-scope fences behind NVLink)
            const int prev = atomicAdd(c.done_cnt + id, 1);
            last = (prev == nblocks - 1);
            if (last) { c.done_cnt[id] = 0; __threadfence(); }   // self-cleaning; acquire the other CTAs' arrivals (their fence + atomic)
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        __syncwarp();                // bar.warp.sync orders memory among the lanes (shfl alone does not): lane 0's acquire of the
                                     // other CTAs' arrivals happens-before the release stores of lanes 1..T-1
        if (last && lane < c.size) {
            // one lane per peer: the T release stores (each a fence + an NVLink round trip) go out side by side, not in sequence;
            // st.release.sys is cumulative over everything ordered before it in this lane, i.e. (through the warp barrier) lane 0's acquire
            const int val = ld_relaxed_gpu(c.step) + 1;
            st_release_sys(tp_flag_ptr(c, lane, id, c.rank), val);
        }
    }
}

// Consumer side: returns when every rank has published flag `id` for the current step.  Call from ONE warp (all
// 32 lanes), then synchronise the CTA.
__device__ __forceinline__ void tp_wait_flags_warp(const TpComm& c, int id) {
    if (c.size <= 1) return;
    const int lane = threadIdx.x & 31;
    const int want = ld_relaxed_gpu(c.step) + 1;
    if (lane < c.size) {
        const int* f = tp_flag_ptr(c, c.rank, id, lane);
        if (ld_acquire_sys(f) < want) {
            const unsigned long long t0 = spin_now();
            while (ld_acquire_sys(f) < want) {
                __nanosleep(20);
                if (spin_now() - t0 > WF_SPIN_BUDGET_NS) {   // peer dead / step counters out of sync: report, do not hang
                    if (c.err) atomicExch(c.err, 1000 + id);
                    break;
                }
            }
        }
    }
    __syncwarp();                    // ld.acquire.sys + the caller's CTA barrier order everything that follows after the peers' writes
}

}  // namespace wf
