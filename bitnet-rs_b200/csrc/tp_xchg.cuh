// tp_xchg.cuh -- cross-GPU exchange inside a kernel (tensor-parallel decode): instead of a NCCL
// collective between two kernels, the kernel that CONSUMES the exchanged data signals its peers, waits for theirs and
// pulls their buffers over NVLink (P2P-mapped symmetric memory) itself, then carries on with the fused epilogue.
//
// Protocol (one `slot` per exchange point of a layer; every rank runs the same sequence of exchanges):
//   * epoch e = counters[slot] + 1, identical on all ranks (they count the same launches);
//   * signal: the rank's data of this exchange was written by the PREVIOUS kernel on the stream (complete and
//     visible); CTA 0 stores e into flags[slot][my rank] of every peer (st.release.sys after a system fence);
//   * wait: every CTA polls its OWN flag array until all ranks show >= e (ld.acquire.sys, bounded spin -> trap);
//   * pull: peer data is read with ld.relaxed.sys (no stale L1 line can exist: first touch in this launch);
//   * the last CTA to finish publishes counters[slot] = e for the next launch (graph replays need no host state).
// Buffer reuse: a rank can only produce the data of exchange n + 2 on the same slot after it has completed exchange
// n + 1, which needed every peer's signal for n + 1, which a peer sends only after its exchange-n kernel has finished
// reading. `phase` splits signal (bit 0) and wait + work (bit 1) for the single-GPU emulation of several ranks, where
// a kernel must never wait for a kernel that has not been launched yet; a real rank passes 3.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int TP_MAX_RANKS = 8, TP_SLOTS = 4;

__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int4 ld_sys_s32x4(const void* p) {
    int4 v;
    asm volatile("ld.relaxed.sys.global.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_sys_s32(const void* p) {
    int v;
    asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_sys_f32(const void* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// all threads of the CTA call it; returns the epoch of this exchange
__device__ __forceinline__ uint32_t xchg_begin(const bitnet_b200_tp_xchg& x) {
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(x.counters + x.slot) + 1u;
    if ((x.phase & 1) && blockIdx.x == 0 && (int)threadIdx.x < x.n_ranks) {
        __threadfence_system();
        st_release_sys_u32(x.peer_flags[threadIdx.x] + x.slot * TP_MAX_RANKS + x.rank, epoch);
    }
    if (x.phase & 2) {
        if ((int)threadIdx.x < x.n_ranks) {
            const uint32_t* f = x.peer_flags[x.rank] + x.slot * TP_MAX_RANKS + threadIdx.x;
            // epochs only grow; signed distance tolerates the 32-bit wrap
            for (uint32_t it = 0; (int32_t)(ld_acquire_sys_u32(f) - epoch) < 0; ++it)
                if (it > (1u << 26)) __trap();  // a lost peer: fail loudly instead of hanging the GPU
        }
        __syncthreads();
    }
    return epoch;
}
__device__ __forceinline__ void xchg_end(const bitnet_b200_tp_xchg& x, uint32_t epoch) {
    if (!(x.phase & 2)) return;
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(x.counters + TP_SLOTS + x.slot, 1u);
        if (t == gridDim.x - 1) {
            x.counters[TP_SLOTS + x.slot] = 0u;
            __threadfence();
            *reinterpret_cast<volatile uint32_t*>(x.counters + x.slot) = epoch;
        }
    }
}

}  // namespace b200
