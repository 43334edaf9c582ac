// Shape-generic fp32 kernels: local training and the scoring forward for ANY autoencoder width.
//
// The reference's models take `hidden_neus` and `latent_dim` as constructor arguments
// (src/Model/Shrink_Autoencoder.py:125-127, src/Model/AutoEncoder.py:125-126) and the thesis runs batch 64;
// the specialised kernels (train_kernel.cu, score_*.cu) are laid out for 115-27-7 and batches <= 32 and keep
// everything in shared memory.  These kernels take D (<= 128, the shard pitch), H, Z and the batch size as
// runtime arguments: same math, same update order (client_trainer.py:360-419), one CTA per client, with
// weights, Adam moments and activations in global memory (L2-resident at these sizes) -- the general path
// behind the fast one, used when the model is not 115-27-7 or the batch exceeds 32 rows.
//
// Packed row layout (layout.py): W1[H][D] b1[H] W2[Z][H] b2[Z] W3[H][Z] b3[H] W4[D][H] b4[D], padded to 32 floats.
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {
namespace gen {

constexpr int GNT = 256;

struct Lay {
    int D, H, Z, oW1, ob1, oW2, ob2, oW3, ob3, oW4, ob4, NP, PP;
};
__host__ __device__ inline Lay make_lay(int D, int H, int Z) {
    Lay L;
    L.D = D; L.H = H; L.Z = Z;
    L.oW1 = 0; L.ob1 = L.oW1 + H * D; L.oW2 = L.ob1 + H; L.ob2 = L.oW2 + Z * H; L.oW3 = L.ob2 + Z; L.ob3 = L.oW3 + H * Z;
    L.oW4 = L.ob3 + H; L.ob4 = L.oW4 + D * H; L.NP = L.ob4 + D; L.PP = (L.NP + 31) / 32 * 32;
    return L;
}

// out[b][n] = act(bias[n] + sum_k in[b][k] * W[n][k]) for b < nb; one (b, n) item per thread
template <bool RELU>
__device__ __forceinline__ void dense(const float* __restrict__ in, int in_pitch, const float* __restrict__ W,
                                      const float* __restrict__ bias, float* __restrict__ out, int nb, int N, int K) {
    for (int it = threadIdx.x; it < nb * N; it += GNT) {
        const int b = it / N, n = it - b * N;
        const float* x = in + (size_t)b * in_pitch;
        const float* w = W + (size_t)n * K;
        float acc = bias[n];
        for (int k = 0; k < K; ++k) acc = fmaf(x[k], w[k], acc);
        out[(size_t)b * N + n] = RELU ? fmaxf(acc, 0.f) : acc;
    }
}

// forward of nb rows; returns (block-wide) sum of squared errors and sum of ||z||; leaves h1, z, h2 and, when
// dX != NULL, dX[b][d] = gscale * (xhat - x) in scratch
__device__ void forward_rows(const Lay& L, const float* __restrict__ w, const float* __restrict__ x, int nb,
                             float* h1, float* z, float* h2, float* dX, float gscale, double* red, double& sse, double& znsum,
                             float* znorm) {
    dense<true>(x, XP, w + L.oW1, w + L.ob1, h1, nb, L.H, L.D);
    __syncthreads();
    dense<false>(h1, L.H, w + L.oW2, w + L.ob2, z, nb, L.Z, L.H);
    __syncthreads();
    dense<true>(z, L.Z, w + L.oW3, w + L.ob3, h2, nb, L.H, L.Z);
    double zn = 0.0;
    for (int b = threadIdx.x; b < nb; b += GNT) {
        float q = 0.f;
        for (int i = 0; i < L.Z; ++i) q = fmaf(z[(size_t)b * L.Z + i], z[(size_t)b * L.Z + i], q);
        const float nrm = sqrtf(q);
        if (znorm) znorm[b] = nrm;
        zn += (double)nrm;
    }
    __syncthreads();
    double s = 0.0;
    for (int it = threadIdx.x; it < nb * L.D; it += GNT) {
        const int b = it / L.D, d = it - b * L.D;
        const float* hh = h2 + (size_t)b * L.H;
        const float* ww = w + L.oW4 + (size_t)d * L.H;
        float acc = w[L.ob4 + d];
        for (int j = 0; j < L.H; ++j) acc = fmaf(hh[j], ww[j], acc);
        const float e = acc - x[(size_t)b * XP + d];
        s += (double)e * (double)e;
        if (dX) dX[(size_t)b * L.D + d] = gscale * e;
    }
    sse = block_sum<double>(s, red);
    znsum = block_sum<double>(zn, red);
}

__device__ __forceinline__ void adam_update(float& w, float& m, float& v, float g, float lr_bc1, float rsq_bc2) {
    m = m + (g - m) * 0.1f;                                   // lerp(m, g, 1 - beta1)
    v = fmaf(v, 0.999f, 0.001f * g * g);
    const float denom = sqrtf(v) * rsq_bc2 + 1e-8f;
    w -= lr_bc1 * (m / denom);
}

__global__ void __launch_bounds__(GNT) train_generic_kernel(const FmTrainGenericArgs a) {
    __shared__ double red[34];
    __shared__ float sc[2];
    const Lay L = make_lay(a.dim, a.hidden, a.latent);
    const int tid = threadIdx.x;
    const int c = a.sel_local[blockIdx.x];
    const int g = a.first_client + c;
    float* w = a.params + (size_t)c * L.PP;
    float* am = a.adam_m + (size_t)c * L.PP;
    float* av = a.adam_v + (size_t)c * L.PP;
    const float* anc = a.anchor + (size_t)c * L.PP;
    float* best = a.best + (size_t)c * L.PP;
    const bool prox = a.fedprox_mu != 0.f;
    const int B = a.batch;
    // scratch of this CTA: h1[B][H] z[B][Z] h2[B][H] dX[B][D] dp2[B][H] dz[B][Z] dp1[B][H] znorm[B]
    float* sp = a.scratch + (size_t)blockIdx.x * a.scratch_stride;
    float* h1 = sp; sp += (size_t)B * L.H;
    float* z = sp; sp += (size_t)B * L.Z;
    float* h2 = sp; sp += (size_t)B * L.H;
    float* dX = sp; sp += (size_t)B * L.D;
    float* dp2 = sp; sp += (size_t)B * L.H;
    float* dz = sp; sp += (size_t)B * L.Z;
    float* dp1 = sp; sp += (size_t)B * L.H;
    float* znorm = sp;
    const float* xt = a.x_train + (size_t)a.train_off[c] * XP;
    const int nt = a.train_off[c + 1] - a.train_off[c];
    const float* xv = a.x_valid + (size_t)a.valid_off[g] * XP;
    const int nv = a.valid_off[g + 1] - a.valid_off[g];
    const int nbt = (nt + B - 1) / B, nbv = (nv + B - 1) / B;
    int t = a.adam_t[c];
    float min_valid = __int_as_float(0x7f800000);
    int worse = 0, epochs_done = 0;
    for (int p = tid; p < L.NP; p += GNT) best[p] = w[p];
    auto prox_term = [&]() -> float {
        if (!prox) return 0.f;
        double pa = 0.0;
        for (int p = tid; p < L.NP; p += GNT) { const double dl = (double)w[p] - (double)anc[p]; pa += dl * dl; }
        return a.fedprox_mu * (float)block_sum<double>(pa, red);
    };
    for (int e = 0; e < a.epochs; ++e) {
        double loss_acc = 0.0;
        for (int s0 = 0; s0 < nt; s0 += B) {
            const int nb = min(B, nt - s0);
            const float* x = xt + (size_t)s0 * XP;
            const float pt = prox_term();                     // proximal term of the loss at the pre-step weights
            double sse, zn;
            forward_rows(L, w, x, nb, h1, z, h2, dX, 2.f / ((float)nb * (float)L.D), red, sse, zn, znorm);
            loss_acc += sse / ((double)nb * L.D) + (double)a.shrink_lambda * zn / nb + (double)pt;
            __syncthreads();
            // ---- backward through the activations ----
            for (int it = tid; it < nb * L.H; it += GNT) {    // dp2 = (dX W4) * (h2 > 0)
                const int b = it / L.H, j = it - b * L.H;
                float acc = 0.f;
                for (int d = 0; d < L.D; ++d) acc = fmaf(dX[(size_t)b * L.D + d], w[L.oW4 + (size_t)d * L.H + j], acc);
                dp2[it] = h2[it] > 0.f ? acc : 0.f;
            }
            __syncthreads();
            for (int it = tid; it < nb * L.Z; it += GNT) {    // dz = dp2 W3 + lambda / nb * z / ||z||
                const int b = it / L.Z, i = it - b * L.Z;
                float acc = 0.f;
                for (int j = 0; j < L.H; ++j) acc = fmaf(dp2[(size_t)b * L.H + j], w[L.oW3 + (size_t)j * L.Z + i], acc);
                const float nrm = znorm[b];
                if (a.shrink_lambda != 0.f && nrm > 0.f) acc += a.shrink_lambda / (float)nb * z[it] / nrm;
                dz[it] = acc;
            }
            __syncthreads();
            for (int it = tid; it < nb * L.H; it += GNT) {    // dp1 = (dz W2) * (h1 > 0)
                const int b = it / L.H, j = it - b * L.H;
                float acc = 0.f;
                for (int i = 0; i < L.Z; ++i) acc = fmaf(dz[(size_t)b * L.Z + i], w[L.oW2 + (size_t)i * L.H + j], acc);
                dp1[it] = h1[it] > 0.f ? acc : 0.f;
            }
            __syncthreads();
            // ---- parameter gradients straight into Adam (torch.optim.Adam defaults, client_trainer.py:66) ----
            ++t;
            if (tid == 0) {
                const double bc1 = 1.0 - pow(0.9, (double)t), bc2 = 1.0 - pow(0.999, (double)t);
                sc[0] = (float)((double)a.lr / bc1);
                sc[1] = (float)(1.0 / sqrt(bc2));
            }
            __syncthreads();
            const float lr_bc1 = sc[0], rsq = sc[1];
            for (int p = tid; p < L.NP; p += GNT) {
                float gsum = 0.f;
                if (p < L.ob1) { const int j = p / L.D, k = p - j * L.D; for (int b = 0; b < nb; ++b) gsum = fmaf(dp1[(size_t)b * L.H + j], x[(size_t)b * XP + k], gsum); }
                else if (p < L.oW2) { const int j = p - L.ob1; for (int b = 0; b < nb; ++b) gsum += dp1[(size_t)b * L.H + j]; }
                else if (p < L.ob2) { const int q = p - L.oW2, i = q / L.H, j = q - i * L.H; for (int b = 0; b < nb; ++b) gsum = fmaf(dz[(size_t)b * L.Z + i], h1[(size_t)b * L.H + j], gsum); }
                else if (p < L.oW3) { const int i = p - L.ob2; for (int b = 0; b < nb; ++b) gsum += dz[(size_t)b * L.Z + i]; }
                else if (p < L.ob3) { const int q = p - L.oW3, j = q / L.Z, i = q - j * L.Z; for (int b = 0; b < nb; ++b) gsum = fmaf(dp2[(size_t)b * L.H + j], z[(size_t)b * L.Z + i], gsum); }
                else if (p < L.oW4) { const int j = p - L.ob3; for (int b = 0; b < nb; ++b) gsum += dp2[(size_t)b * L.H + j]; }
                else if (p < L.ob4) { const int q = p - L.oW4, d = q / L.H, j = q - d * L.H; for (int b = 0; b < nb; ++b) gsum = fmaf(dX[(size_t)b * L.D + d], h2[(size_t)b * L.H + j], gsum); }
                else { const int d = p - L.ob4; for (int b = 0; b < nb; ++b) gsum += dX[(size_t)b * L.D + d]; }
                float wv = w[p], mv = am[p], vv = av[p];
                if (prox) gsum = fmaf(2.f * a.fedprox_mu, wv - anc[p], gsum);
                adam_update(wv, mv, vv, gsum, lr_bc1, rsq);
                w[p] = wv; am[p] = mv; av[p] = vv;
            }
            __syncthreads();
        }
        // ---- validation pass, epoch bookkeeping (client_trainer.py:385-417) ----
        double vloss = 0.0;
        const float ptv = prox_term();
        for (int s0 = 0; s0 < nv; s0 += B) {
            const int nb = min(B, nv - s0);
            double sse, zn;
            forward_rows(L, w, xv + (size_t)s0 * XP, nb, h1, z, h2, nullptr, 0.f, red, sse, zn, nullptr);
            vloss += sse / ((double)nb * L.D) + (double)a.shrink_lambda * zn / nb + (double)ptv;
            __syncthreads();
        }
        const float train_loss = (float)(loss_acc / (double)max(nbt, 1));
        const float valid_loss = nbv > 0 ? (float)(vloss / (double)nbv) : __int_as_float(0x7fc00000);
        if (tid == 0 && e < a.log_stride) {
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 0] = train_loss;
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 1] = valid_loss;
        }
        epochs_done = e + 1;
        if (nbv == 0) continue;                               // run(train_loader, None): all epochs, nothing saved
        if (valid_loss < min_valid) {
            min_valid = valid_loss; worse = 0;
            for (int p = tid; p < L.NP; p += GNT) best[p] = w[p];
        } else if (++worse >= a.patience) break;
    }
    if (tid == 0) {
        a.adam_t[c] = t;
        a.epochs_run[c] = epochs_done;
        a.trained_rounds[c] += 1;
        for (int e = epochs_done; e < a.log_stride; ++e) {
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 0] = 0.f;
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 1] = 0.f;
        }
    }
}

// scoring forward: one CTA per (model slot, tile of up to 16 rows); writes the per-row SSE and / or the latent
constexpr int GT = 16;
__global__ void __launch_bounds__(GNT) score_generic_kernel(const FmScoreGenericArgs a) {
    extern __shared__ __align__(16) float gsm[];
    const Lay L = make_lay(a.dim, a.hidden, a.latent);
    const int* tl = a.tiles + 4 * (size_t)blockIdx.x;
    const int slot = tl[0], row0 = tl[1], nrows = tl[2], out0 = tl[3];
    const int cg = a.slot_client ? a.slot_client[slot] : a.first_client + slot;
    const int c = cg - a.first_client;
    if (c < 0 || c >= a.n_local) return;                      // owned by another rank
    const float* w = a.slab + (size_t)c * L.PP;
    const float* x = a.X + (size_t)row0 * XP;
    float* h1 = gsm;
    float* z = h1 + GT * L.H;
    float* h2 = z + GT * L.Z;
    dense<true>(x, XP, w + L.oW1, w + L.ob1, h1, nrows, L.H, L.D);
    __syncthreads();
    dense<false>(h1, L.H, w + L.oW2, w + L.ob2, z, nrows, L.Z, L.H);
    __syncthreads();
    if (a.out_lat)
        for (int it = threadIdx.x; it < nrows * L.Z; it += GNT) {
            const int b = it / L.Z, i = it - b * L.Z;
            a.out_lat[(size_t)(out0 + b) * a.lat_pitch + i] = z[it];
        }
    if (!a.out_sse) return;
    dense<true>(z, L.Z, w + L.oW3, w + L.ob3, h2, nrows, L.H, L.Z);
    __syncthreads();
    // one warp per row: lanes stride the output features
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    for (int b = wp; b < nrows; b += GNT / 32) {
        float s = 0.f;
        for (int d = lane; d < L.D; d += 32) {
            const float* ww = w + L.oW4 + (size_t)d * L.H;
            float acc = w[L.ob4 + d];
            for (int j = 0; j < L.H; ++j) acc = fmaf(h2[b * L.H + j], ww[j], acc);
            const float e = acc - x[(size_t)b * XP + d];
            s = fmaf(e, e, s);
        }
        s = warp_sum(s);
        if (lane == 0) a.out_sse[out0 + b] = s;
    }
}

}  // namespace gen
}  // namespace fm

extern "C" int64_t fm_train_generic_scratch_floats(int32_t dim, int32_t hidden, int32_t latent, int32_t batch) {
    return (int64_t)batch * (4 * (int64_t)hidden + 2 * (int64_t)latent + dim + 1) + 32;   // h1 h2 dp2 dp1 | z dz | dX | znorm
}

extern "C" int fm_train_generic(const FmTrainGenericArgs* a, void* stream) {
    using namespace fm::gen;
    if (a->n_sel <= 0) return 0;
    if (a->dim < 1 || a->dim > fm::XP || a->hidden < 1 || a->latent < 1 || a->batch < 1 || a->batch > 4096)
        return (int)cudaErrorInvalidValue;
    if (a->scratch == nullptr || a->scratch_stride < fm_train_generic_scratch_floats(a->dim, a->hidden, a->latent, a->batch))
        return (int)cudaErrorInvalidValue;
    train_generic_kernel<<<a->n_sel, GNT, 0, (cudaStream_t)stream>>>(*a);
    return (int)cudaGetLastError();
}

extern "C" int fm_score_generic(const FmScoreGenericArgs* a, void* stream) {
    using namespace fm::gen;
    if (a->n_tiles <= 0) return 0;
    if (a->dim < 1 || a->dim > fm::XP || a->hidden < 1 || a->latent < 1) return (int)cudaErrorInvalidValue;
    const size_t smem = (size_t)GT * (2 * (size_t)a->hidden + a->latent) * sizeof(float);
    if (smem > 200 * 1024) return (int)cudaErrorInvalidValue;          // hidden width beyond ~1,500: not served here
    cudaError_t err = cudaFuncSetAttribute(score_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    score_generic_kernel<<<a->n_tiles, GNT, smem, (cudaStream_t)stream>>>(*a);
    return (int)cudaGetLastError();
}
