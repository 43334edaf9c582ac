// Shared device code of the fp32 row-scoring forward (used by score_kernel.cu and gossip_kernel.cu).
#pragma once
#include "common.cuh"

namespace fm {

constexpr int SNT = 256;           // 8 warps
constexpr int TILE = 64;           // rows per tile, 8 per warp
constexpr int W1TP = 33;           // W1T[k][j], odd pitch: conflict-free transposed fill
constexpr int W4TP = 129;          // W4T[j][d]
constexpr int K1 = 116, J4 = 28;

struct ScoreSmem {
    alignas(16) float x[2][TILE * XP];
    float W1T[K1 * W1TP];
    float W4T[J4 * W4TP];
    float W2T[32 * 8];             // [j][i]
    float W3T[8 * 32];             // [i][j]
    float b1[32], b2[8], b3[32], b4[128];
    alignas(16) float h1s[8][4][32];
    alignas(16) float zs[8][4][8];
    alignas(16) float h2s[8][4][32];
    alignas(8) uint64_t bar[2];
};

__device__ __forceinline__ void load_model_transposed(ScoreSmem& S, const float* __restrict__ row) {
    const int tid = threadIdx.x;
    for (int i = tid; i < K1 * W1TP; i += SNT) S.W1T[i] = 0.f;
    for (int i = tid; i < J4 * W4TP; i += SNT) S.W4T[i] = 0.f;
    for (int i = tid; i < 256; i += SNT) { S.W2T[i] = 0.f; S.W3T[i] = 0.f; }
    if (tid < 32) { S.b1[tid] = 0.f; S.b3[tid] = 0.f; }
    if (tid < 8) S.b2[tid] = 0.f;
    if (tid < 128) S.b4[tid] = 0.f;
    __syncthreads();
    for (int p = tid; p < NP; p += SNT) {
        const float v = row[p];
        if (p < OFF_B1) { const int j = p / D, k = p - j * D; S.W1T[k * W1TP + j] = v; }
        else if (p < OFF_W2) S.b1[p - OFF_B1] = v;
        else if (p < OFF_B2) { const int q = p - OFF_W2, i = q / H, j = q - i * H; S.W2T[j * 8 + i] = v; }
        else if (p < OFF_W3) S.b2[p - OFF_B2] = v;
        else if (p < OFF_B3) { const int q = p - OFF_W3, j = q / Z, i = q - j * Z; S.W3T[i * 32 + j] = v; }
        else if (p < OFF_W4) S.b3[p - OFF_B3] = v;
        else if (p < OFF_B4) { const int q = p - OFF_W4, d = q / H, j = q - d * H; S.W4T[j * W4TP + d] = v; }
        else S.b4[p - OFF_B4] = v;
    }
    __syncthreads();
}

// One warp scores 4 rows (xr = first row in shared memory).  Returns per-row SSE in sse[4] (all lanes).
__device__ __forceinline__ void score4(ScoreSmem& S, const float* __restrict__ xr, int w, int lane, bool need_sse,
                                       float sse[4], float zout[4]) {
    // ---- layer 1: lane = hidden unit j ----
    float a0 = S.b1[lane], a1 = a0, a2 = a0, a3 = a0;
#pragma unroll 4
    for (int k = 0; k < K1; k += 4) {
        const float4 x0 = *reinterpret_cast<const float4*>(xr + k);
        const float4 x1 = *reinterpret_cast<const float4*>(xr + XP + k);
        const float4 x2 = *reinterpret_cast<const float4*>(xr + 2 * XP + k);
        const float4 x3 = *reinterpret_cast<const float4*>(xr + 3 * XP + k);
        const float w0 = S.W1T[k * W1TP + lane], w1 = S.W1T[(k + 1) * W1TP + lane];
        const float w2 = S.W1T[(k + 2) * W1TP + lane], w3 = S.W1T[(k + 3) * W1TP + lane];
        a0 = fmaf(x0.x, w0, fmaf(x0.y, w1, fmaf(x0.z, w2, fmaf(x0.w, w3, a0))));
        a1 = fmaf(x1.x, w0, fmaf(x1.y, w1, fmaf(x1.z, w2, fmaf(x1.w, w3, a1))));
        a2 = fmaf(x2.x, w0, fmaf(x2.y, w1, fmaf(x2.z, w2, fmaf(x2.w, w3, a2))));
        a3 = fmaf(x3.x, w0, fmaf(x3.y, w1, fmaf(x3.z, w2, fmaf(x3.w, w3, a3))));
    }
    S.h1s[w][0][lane] = fmaxf(a0, 0.f); S.h1s[w][1][lane] = fmaxf(a1, 0.f);
    S.h1s[w][2][lane] = fmaxf(a2, 0.f); S.h1s[w][3][lane] = fmaxf(a3, 0.f);
    __syncwarp();
    // ---- layer 2: lane = (row r, latent i), 28 active lanes ----
    {
        const int r = lane / Z < 4 ? lane / Z : 3, i = lane - (lane / Z) * Z;
        float z = S.b2[i];
#pragma unroll
        for (int j = 0; j < H; ++j) z = fmaf(S.h1s[w][r][j], S.W2T[j * 8 + i], z);
        if (lane < 4 * Z) S.zs[w][r][i] = z;
        zout[0] = z;  // lane-private: latent (r, i) for lanes < 28
    }
    __syncwarp();
    if (!need_sse) { sse[0] = sse[1] = sse[2] = sse[3] = 0.f; return; }
    // ---- layer 3: lane = hidden unit j ----
    {
        float c0 = S.b3[lane], c1 = c0, c2 = c0, c3 = c0;
#pragma unroll
        for (int i = 0; i < Z; ++i) {
            const float wv = S.W3T[i * 32 + lane];
            c0 = fmaf(S.zs[w][0][i], wv, c0); c1 = fmaf(S.zs[w][1][i], wv, c1);
            c2 = fmaf(S.zs[w][2][i], wv, c2); c3 = fmaf(S.zs[w][3][i], wv, c3);
        }
        S.h2s[w][0][lane] = fmaxf(c0, 0.f); S.h2s[w][1][lane] = fmaxf(c1, 0.f);
        S.h2s[w][2][lane] = fmaxf(c2, 0.f); S.h2s[w][3][lane] = fmaxf(c3, 0.f);
    }
    __syncwarp();
    // ---- layer 4 + squared error: lane = output feature d (4 passes of 32) ----
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int dp = 0; dp < 4; ++dp) {
        const int d = dp * 32 + lane;
        float o0 = S.b4[d], o1 = o0, o2 = o0, o3 = o0;
#pragma unroll
        for (int j = 0; j < J4; j += 4) {
            const float4 h0 = *reinterpret_cast<const float4*>(&S.h2s[w][0][j]);
            const float4 h1 = *reinterpret_cast<const float4*>(&S.h2s[w][1][j]);
            const float4 h2 = *reinterpret_cast<const float4*>(&S.h2s[w][2][j]);
            const float4 h3 = *reinterpret_cast<const float4*>(&S.h2s[w][3][j]);
            const float w0 = S.W4T[j * W4TP + d], w1 = S.W4T[(j + 1) * W4TP + d];
            const float w2 = S.W4T[(j + 2) * W4TP + d], w3 = S.W4T[(j + 3) * W4TP + d];
            o0 = fmaf(h0.x, w0, fmaf(h0.y, w1, fmaf(h0.z, w2, fmaf(h0.w, w3, o0))));
            o1 = fmaf(h1.x, w0, fmaf(h1.y, w1, fmaf(h1.z, w2, fmaf(h1.w, w3, o1))));
            o2 = fmaf(h2.x, w0, fmaf(h2.y, w1, fmaf(h2.z, w2, fmaf(h2.w, w3, o2))));
            o3 = fmaf(h3.x, w0, fmaf(h3.y, w1, fmaf(h3.z, w2, fmaf(h3.w, w3, o3))));
        }
        if (d < D) {
            const float e0 = o0 - xr[d], e1 = o1 - xr[XP + d], e2 = o2 - xr[2 * XP + d], e3 = o3 - xr[3 * XP + d];
            s0 = fmaf(e0, e0, s0); s1 = fmaf(e1, e1, s1); s2 = fmaf(e2, e2, s2); s3 = fmaf(e3, e3, s3);
        }
    }
    sse[0] = warp_sum(s0); sse[1] = warp_sum(s1); sse[2] = warp_sum(s2); sse[3] = warp_sum(s3);
    __syncwarp();
}

}  // namespace fm
