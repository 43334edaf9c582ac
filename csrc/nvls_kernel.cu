// NVLS (NVLink SHARP) paths: reductions done INSIDE the NVSwitch through a multicast mapping of a symmetric buffer
// (multimem.ld_reduce) and broadcasts written once and replicated by the switch (multimem.st).
//
// Reference behaviour being replaced: the aggregator gathers every selected state_dict and averages them
// (src/Trainer/client_trainer.py:107-130, :306-335) and hands the result to every client (src/main.py:298-300).
//
//  * fm_merge_nvls -- the elected-mode merge without pulling rows across GPUs.  The MSE board is replicated on
//    every rank, so every rank knows ALL the normalised weights; it therefore pre-scales and sums only the selected
//    rows it OWNS into a symmetric partial buffer (local reads), and the switch adds the partials:
//        phase 1  partial_r = sum over owned selected k of w_k * theta_k          (local memory only)
//        phase 2  theta[slice_r] = multimem.ld_reduce.add(partial[slice_r])        (one read, reduced in the switch)
//                 multimem.st(theta[slice_r])                                      (one write, lands on every GPU)
//    Each GPU moves 2 x 27 KB / N over NVLink instead of pulling n_sel x 27 KB (13.5 MB at 1,000 clients), every rank
//    receives bit-identical theta (each element is reduced exactly once), and the same arrival counter that used to
//    mean "done pulling" now means "all slices of theta have landed".
//  * fm_nvls_allreduce_slice -- phase 2 of the wide-model all-reduce (csrc/allreduce_p2p.cu) in the same way: the
//    reduce-scatter + all-gather by pulling becomes one ld_reduce + one st per element.
//
// The multicast pointers come from torch.distributed._symmetric_memory (parallel/nvls.py does the rendezvous); hosts
// without multicast support keep the peer-load kernels.  Loads and stores here are multimem.* on a multicast address
// or plain accesses to this GPU's own memory -- no TMA is involved.
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void nv_signal(unsigned* addr) {
    asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(addr) : "memory");
}
__device__ __forceinline__ unsigned nv_load(const unsigned* addr) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void nv_wait(const unsigned* arrive, unsigned expect) {
    if (threadIdx.x == 0) while ((int)(nv_load(arrive) - expect) < 0) __nanosleep(64);
    __syncthreads();
}
__device__ __forceinline__ unsigned* nv_arrive(float* ctl, int which) {
    return reinterpret_cast<unsigned*>(ctl + 2 * FM_MAX_SEL) + which;
}
// every thread fences its writes; the last block of the grid bumps arrive[which] on every rank
__device__ __forceinline__ void nv_signal_all(float* const* peer_ctl, int world, int which, unsigned* grid_counter) {
    __shared__ unsigned s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(grid_counter, 1u);
    __syncthreads();
    if (s_last == gridDim.x - 1) {
        if (threadIdx.x == 0) *grid_counter = 0u;
        if ((int)threadIdx.x < world) nv_signal(nv_arrive(peer_ctl[threadIdx.x], which));
    }
}
__device__ __forceinline__ void nv_owner_of(int g, int n_global, int world, int& rank, int& local) {
    const int base = n_global / world, rem = n_global - base * world;
    const int cut = rem * (base + 1);
    if (g < cut) { rank = g / (base + 1); local = g - rank * (base + 1); }
    else { const int q = (g - cut) / base; rank = rem + q; local = (g - cut) - q * base; }
}

constexpr int NV_NT = 256;
constexpr int NV_P4 = P_PAD / 4;                      // 1,696 float4 per model

__global__ void __launch_bounds__(NV_NT) merge_nvls_kernel(const FmMergeNvlsArgs a) {
    extern __shared__ __align__(16) unsigned char nsm[];
    float* wts = reinterpret_cast<float*>(nsm);                                   // [n_sel]
    int* own = reinterpret_cast<int*>(nsm + (size_t)a.n_sel * sizeof(float));    // [n_sel] local row or -1
    __shared__ double red[34];
    const int tid = threadIdx.x, n = a.n_sel, ap = a.agg_pos[0];
    if (ap < 0) {                         // nothing to merge: keep both counters in step with the host's sequence numbers
        if (blockIdx.x == 0 && tid < a.world) {
            nv_signal(nv_arrive(a.peer_ctl[tid], FM_ARRIVE_PARTIAL));
            nv_signal(nv_arrive(a.peer_ctl[tid], FM_ARRIVE_DONE));
        }
        return;
    }
    if (a.arrive_mse != nullptr) nv_wait(a.arrive_mse, a.expect_mse);            // the MSE board of this rank is complete
    // ---- weights: same arithmetic as merge_kernel (every CTA recomputes them) ----
    if (a.mode == 0) {
        for (int k = tid; k < n; k += NV_NT) wts[k] = 1.f / (float)n;
    } else {
        double part = 0.0;
        for (int k = tid; k < n; k += NV_NT) {
            int src = k;
            if (a.compat && n > 1 && k == ap && ap > 0) src = ap - 1;
            const float s = 1.f / a.mse[src];
            wts[k] = s;
            part += (double)s;
        }
        const double tot = block_sum<double>(part, red);
        for (int k = tid; k < n; k += NV_NT) wts[k] = (float)((double)wts[k] / tot);
        __syncthreads();
        if (a.compat && n > 1 && tid == 0) {
            const int last = (ap != n - 1) ? n - 1 : n - 2;
            const float moved = wts[ap];
            wts[ap] = 0.f;
            wts[last] += moved;
        }
    }
    for (int k = tid; k < n; k += NV_NT) {
        int r, l;
        nv_owner_of(a.sel[k], a.n_global, a.world, r, l);
        own[k] = r == a.rank ? l : -1;
    }
    __syncthreads();
    if (a.weights_out && blockIdx.x == 0)
        for (int k = tid; k < n; k += NV_NT) a.weights_out[k] = wts[k];
    // ---- phase 1: this rank's share of the weighted sum, selection order, local rows only ----
    const int i = blockIdx.x * NV_NT + tid;
    if (i < NV_P4) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = 0; k < n; ++k) {
            const int l = own[k];
            if (l < 0) continue;
            const float4 v = reinterpret_cast<const float4*>(a.slab + (size_t)l * P_PAD)[i];
            const float w = wts[k];
            acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y); acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
        }
        reinterpret_cast<float4*>(a.partial_local)[i] = acc;
    }
    nv_signal_all(a.peer_ctl, a.world, FM_ARRIVE_PARTIAL, a.grid_counter);
    nv_wait(a.arrive_partial, a.expect_partial);                                   // every rank's partial is in place
    // ---- phase 2: my slice of theta, reduced in the switch, stored to every GPU ----
    const int per = (NV_P4 + a.world - 1) / a.world;
    const int s0 = per * a.rank, s1 = min(NV_P4, s0 + per);
    for (int j = s0 + blockIdx.x * NV_NT + tid; j < s1; j += gridDim.x * NV_NT) {
        const float4 v = multimem_ld_reduce_add(a.mc_partial + 4 * (size_t)j);
        multimem_st(a.mc_theta + 4 * (size_t)j, v);
    }
    if (blockIdx.x == 0 && tid == 0) a.agg_count[a.sel[ap]] += 1;
    nv_signal_all(a.peer_ctl, a.world, FM_ARRIVE_DONE, a.grid_counter + 1);
}

// stream-order a later kernel behind a cross-GPU arrival counter
__global__ void flag_wait_kernel(const unsigned* arrive, unsigned expect) {
    while ((int)(nv_load(arrive) - expect) < 0) __nanosleep(64);
}

// wide models: reduced[slice_r] = sum over ranks of partial[slice_r] through the multicast mapping, written back to
// every rank's buffer in place (the buffer then holds the complete all-reduced vector on every GPU)
__global__ void __launch_bounds__(256) nvls_allreduce_slice_kernel(float* mc, size_t slice0_4, size_t slice_n4,
                                                                   const unsigned* arrive, unsigned expect,
                                                                   float* const* peer_ctl, int world, int which,
                                                                   unsigned* grid_counter) {
    nv_wait(arrive, expect);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < slice_n4; i += (size_t)gridDim.x * blockDim.x) {
        float* p = mc + 4 * (slice0_4 + i);
        multimem_st(p, multimem_ld_reduce_add(p));
    }
    nv_signal_all(peer_ctl, world, which, grid_counter);
}

}  // namespace fm

extern "C" int fm_merge_nvls(const FmMergeNvlsArgs* args, void* stream) {
    using namespace fm;
    if (args->n_sel <= 0 || args->n_sel > FM_MAX_SEL || args->world < 2 || args->world > 32) return (int)cudaErrorInvalidValue;
    if (!args->mc_partial || !args->mc_theta || !args->peer_ctl) return (int)cudaErrorInvalidValue;
    const int grid = (NV_P4 + NV_NT - 1) / NV_NT;              // 7 CTAs: all resident, so the in-kernel waits are safe
    merge_nvls_kernel<<<grid, NV_NT, args->n_sel * (sizeof(float) + sizeof(int)), (cudaStream_t)stream>>>(*args);
    return (int)cudaGetLastError();
}

extern "C" int fm_flag_wait(const uint32_t* arrive, uint32_t expect, void* stream) {
    fm::flag_wait_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(arrive, expect);
    return (int)cudaGetLastError();
}

extern "C" int fm_nvls_allreduce_slice(float* mc, int32_t world, int32_t rank, int64_t n_floats, const uint32_t* arrive,
                                       uint32_t expect, float* const* peer_ctl, int32_t which, uint32_t* grid_counter,
                                       void* stream) {
    if (n_floats % 4 || !mc || !peer_ctl) return (int)cudaErrorInvalidValue;
    const size_t p4 = (size_t)n_floats / 4, per = (p4 + world - 1) / world;
    const size_t s0 = per * rank, sn = s0 >= p4 ? 0 : (p4 - s0 < per ? p4 - s0 : per);
    size_t g = (sn + 255) / 256;
    // in-kernel cross-rank wait: keep the grid within one resident wave
    const int grid = (int)(g < 1 ? 1 : (g > 148 * 4 ? 148 * 4 : g));
    fm::nvls_allreduce_slice_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(mc, s0, sn, arrive, expect, peer_ctl, world, which,
                                                                          grid_counter);
    return (int)cudaGetLastError();
}
