// work_dispenser.cuh -- how the persistent dual-frame kernels hand frame pairs to their warps.
//
// Static striding (pair q, q + grid*warps, ...) gives every warp the same NUMBER of pairs, but SMs do not run
// at the same speed (L2-die distance, two CTAs sharing one SM's L1 pipe), so the slowest SM sets the kernel
// time while the others idle: ncu showed 7.1 of 8 resident warps active on average.  With kDyn the pairs come
// from one device-wide atomic counter instead (+6 ... +10 % on B200, profiles/README.md).  Because the input of
// the NEXT pair is prefetched by TMA while the current one is transformed, a warp always knows two units
// (`base`, `nbase`); a third grab is kept in flight in the grabbing thread (`pend`) so that the atomic's round
// trip to L2 is never waited for.
//
//   kFree = false: the CTA's warps stay in lock step (block barrier at the top of every trip, see PHASE_SYNC in
//                  the kernels); thread 0 grabs kWarps consecutive pairs for the CTA and publishes the base
//                  through shared memory.
//   kFree = true : every warp runs on its own; lane 0 grabs one pair at a time.
// The counter must be zero at launch (the launchers enqueue a 8-byte memset in front of the kernel).
#pragma once
#include <cuda_runtime.h>

namespace r2iq {

template <bool kDyn, bool kFree, int kWarps>
struct PairDispenser {
    long long base, nbase, nnbase, stride;
    unsigned long long pend;
    bool grabber;

    static constexpr unsigned long long kGrab = kFree ? 1 : kWarps;

    // s_base: three long long of shared memory per CTA (used only with kDyn && !kFree)
    __device__ __forceinline__ void init(unsigned long long* counter, long long* s_base, int lane, int warp) {
        stride = (long long)gridDim.x * kWarps;
        grabber = kFree ? (lane == 0) : (threadIdx.x == 0);
        pend = 0;
        if constexpr (kDyn) {
            unsigned long long a = 0, b = 0;
            if (grabber) {
                a = atomicAdd(counter, kGrab);
                b = atomicAdd(counter, kGrab);
                pend = atomicAdd(counter, kGrab);
            }
            if constexpr (kFree) {
                base = (long long)__shfl_sync(0xffffffffu, a, 0);
                nbase = (long long)__shfl_sync(0xffffffffu, b, 0);
            } else {
                if (grabber) {
                    s_base[0] = (long long)a;
                    s_base[1] = (long long)b;
                }
                __syncthreads();
                base = s_base[0];
                nbase = s_base[1];
            }
        } else {
            base = (long long)blockIdx.x * kWarps + (kFree ? warp : 0);
            nbase = base + stride;
        }
        nnbase = nbase + stride;
    }
    // pair of this warp in the current / the next trip (>= n_pairs: nothing to do)
    __device__ __forceinline__ long long cur(int warp) const { return base + (kFree ? 0 : warp); }
    __device__ __forceinline__ long long next(int warp) const { return nbase + (kFree ? 0 : warp); }
    // loop condition: CTA-uniform without kFree
    __device__ __forceinline__ bool more(long long n_pairs) const { return base < n_pairs; }
    // top of a trip: learn the unit after next; contains the block barrier of the lock-step variants
    __device__ __forceinline__ void top(unsigned long long* counter, long long* s_base) {
        nnbase = nbase + stride;
        if constexpr (kDyn && !kFree) {
            if (grabber) {
                s_base[2] = (long long)pend;     // read by everyone after the barrier below, rewritten only after
                pend = atomicAdd(counter, kGrab);   // every thread has passed a later barrier of this trip
            }
        }
        if constexpr (!kFree) __syncthreads();
        if constexpr (kDyn) {
            if constexpr (kFree) {
                const unsigned long long t = pend;
                if (grabber) pend = atomicAdd(counter, kGrab);
                nnbase = (long long)__shfl_sync(0xffffffffu, t, 0);
            } else {
                nnbase = s_base[2];
            }
        }
    }
    __device__ __forceinline__ void advance() {
        base = nbase;
        nbase = nnbase;
    }
};

}  // namespace r2iq
