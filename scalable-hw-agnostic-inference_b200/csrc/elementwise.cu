// HBM-bound kernels of the hot path: GroupNorm(+SiLU, +concat), LayerNorm, im2col for the two convs TMA
// cannot tile, nearest 2x upsample, CFG + scheduler step (K4), the timestep-embedding small-M linears (K5),
// row softmax for the VAE's single-head attention, uint8 post-process (K7), casts.
// All are coalesced, 8/16-byte vectorised, grid sized in multiples of the SM count where the shape allows.
// Element types are selected at run time by a dtype code (warp-uniform branch; these kernels are bandwidth-bound).
// Algorithmic bytes per element are listed per kernel in DESIGN.md section 4.
#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace b200sd {
using namespace ptx;

// x * sigmoid(x) with the approximate reciprocal (MUFU.RCP, ~1 ulp): the IEEE division's range-handling slow path was a
// third of gn_apply's instructions (ncu: issue slots 64 % busy, XU 59 %, in a kernel that should only wait for HBM).
// 1 + exp(-x) overflows to +inf only for x < -88, where the quotient is the correct limit 0.
__device__ __forceinline__ float silu_f(float x) { return __fdividef(x, 1.0f + __expf(-x)); }

// ---- 4-wide loads/stores; idx counts elements ---------------------------------------------------------------
__device__ __forceinline__ float4 load4(const void* base, size_t idx, int dt) {
    if (dt == B200SD_F32) return *reinterpret_cast<const float4*>(static_cast<const float*>(base) + idx);
    const uint2 v = *reinterpret_cast<const uint2*>(static_cast<const uint16_t*>(base) + idx);
    const bool f16 = dt == B200SD_F16;
    return make_float4(h_lo(v.x, f16), h_hi(v.x, f16), h_lo(v.y, f16), h_hi(v.y, f16));
}
__device__ __forceinline__ float load1(const void* base, size_t idx, int dt) {
    if (dt == B200SD_F32) return static_cast<const float*>(base)[idx];
    if (dt == B200SD_F16) return __half2float(static_cast<const __half*>(base)[idx]);
    return __bfloat162float(static_cast<const __nv_bfloat16*>(base)[idx]);
}
__device__ __forceinline__ void store4_h(void* base, size_t idx, float4 v, bool f16) {
    *reinterpret_cast<uint2*>(static_cast<uint16_t*>(base) + idx) =
        make_uint2(pack_h2(v.x, v.y, f16), pack_h2(v.z, v.w, f16));
}
__device__ __forceinline__ void store8_h(void* base, size_t idx, const float (&v)[8], bool f16) {
    *reinterpret_cast<uint4*>(static_cast<uint16_t*>(base) + idx) =
        make_uint4(pack_h2(v[0], v[1], f16), pack_h2(v[2], v[3], f16), pack_h2(v[4], v[5], f16),
                   pack_h2(v[6], v[7], f16));
}

// =============================================================================================== GroupNorm
// Pass 1: per (image, pixel-split) partial sums per group, written (not accumulated) to scratch so the result
// is deterministic.  Pass 2 reduces the splits in a fixed order, then normalises.
// Thread layout: blockDim = vpp * lanes, vpp = C/VEC vectors per pixel; thread -> fixed VEC channels (VEC = 8 when the
// channel counts allow: 16-byte loads of 16-bit inputs and 16-byte stores of the 16-bit outputs).
template <int VEC>
__device__ __forceinline__ void loadv(const void* base, size_t idx, int dt, float (&v)[VEC]) {
    if (VEC == 4) {
        const float4 a = load4(base, idx, dt);
        v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w;
    } else if (dt == B200SD_F32) {
        const float4 a = *reinterpret_cast<const float4*>(static_cast<const float*>(base) + idx);
        const float4 b = *reinterpret_cast<const float4*>(static_cast<const float*>(base) + idx + 4);
        v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
    } else {
        const uint4 t = *reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(base) + idx);
        const bool f16 = dt == B200SD_F16;
        v[0] = h_lo(t.x, f16), v[1] = h_hi(t.x, f16), v[2] = h_lo(t.y, f16), v[3] = h_hi(t.y, f16);
        v[4] = h_lo(t.z, f16), v[5] = h_hi(t.z, f16), v[6] = h_lo(t.w, f16), v[7] = h_hi(t.w, f16);
    }
}
template <int VEC>
__device__ __forceinline__ void storev_h(void* base, size_t idx, const float (&v)[VEC], bool f16) {
    if (VEC == 4) {
        store4_h(base, idx, make_float4(v[0], v[1], v[2], v[3]), f16);
    } else {
        *reinterpret_cast<uint4*>(static_cast<uint16_t*>(base) + idx) =
            make_uint4(pack_h2(v[0], v[1], f16), pack_h2(v[2], v[3], f16), pack_h2(v[4], v[5], f16),
                       pack_h2(v[6], v[7], f16));
    }
}

template <int VEC>
__global__ void gn_stats_kernel(const void* __restrict__ x1, int c1, const void* __restrict__ x2, int c2, int dt,
                                int hw, int groups, int splits, float* __restrict__ partial) {
    pdl_sync();
    extern __shared__ float sm[];  // [lanes][C] sums, then [lanes][C] sums of squares
    const int C = c1 + c2;
    const int vpp = C / VEC;
    const int lanes = blockDim.x / vpp;
    const int pl = threadIdx.x / vpp;  // pixel lane
    const int vec = threadIdx.x - pl * vpp;
    const int ch = vec * VEC;
    const int n = blockIdx.y, split = blockIdx.x;
    const int per = (hw + splits - 1) / splits;
    const int p0 = split * per;
    const int p1 = min(hw, p0 + per);
    const bool second = ch >= c1;
    const void* src = second ? x2 : x1;
    const int cs = second ? c2 : c1;
    const size_t off = (size_t)n * hw * cs + (second ? ch - c1 : ch);
    float s[VEC], q[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) s[j] = 0.f, q[j] = 0.f;
    constexpr int U = VEC == 4 ? 4 : 2;  // independent loads in flight per thread (64 / 32-64 bytes; 4 x 8 measured no better)
    int p = p0 + pl;
    // these tensors are read once, so latency hiding is all that matters
    for (; p + (U - 1) * lanes < p1; p += U * lanes) {
        float v[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u) loadv<VEC>(src, off + (size_t)(p + u * lanes) * cs, dt, v[u]);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < VEC; ++j) s[j] += v[u][j], q[j] += v[u][j] * v[u][j];
    }
    for (; p < p1; p += lanes) {
        float v[VEC];
        loadv<VEC>(src, off + (size_t)p * cs, dt, v);
#pragma unroll
        for (int j = 0; j < VEC; ++j) s[j] += v[j], q[j] += v[j] * v[j];
    }
    float* ss = sm + pl * C + ch;
    float* qq = sm + lanes * C + pl * C + ch;
#pragma unroll
    for (int j = 0; j < VEC; ++j) ss[j] = s[j], qq[j] = q[j];
    __syncthreads();
    const int cpg = C / groups;
    for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        float a = 0.f, b = 0.f;
        for (int l = 0; l < lanes; ++l)
            for (int c = g * cpg; c < (g + 1) * cpg; ++c) a += sm[l * C + c], b += sm[lanes * C + l * C + c];
        float* out = partial + (((size_t)n * splits + split) * groups + g) * 2;
        out[0] = a, out[1] = b;
    }
}

// The element formats and the activation are template parameters: selected at run time per element, every load carried
// both 16-bit decodings plus a select and every store both encodings (ncu on the level-0 shape: 64 % of the issue slots
// busy in a kernel that should only wait for HBM).
template <int VEC, int DT, bool OF16, bool SILU>
__global__ void gn_apply_kernel(const void* __restrict__ x1, int c1, const void* __restrict__ x2, int c2,
                                int hw, int groups, int splits, const float* __restrict__ partial, float eps,
                                const float* __restrict__ gamma, const float* __restrict__ beta,
                                void* __restrict__ y, void* __restrict__ raw, int chunks, float stat_hw) {
    constexpr int dt = DT;
    constexpr bool silu = SILU;
    constexpr int out_f16 = OF16 ? 1 : 0;
    pdl_sync();
    __shared__ float s_mean[64], s_rstd[64];
    const int C = c1 + c2;
    const int vpp = C / VEC;
    const int lanes = blockDim.x / vpp;
    const int pl = threadIdx.x / vpp;
    const int vec = threadIdx.x - pl * vpp;
    const int ch = vec * VEC;
    const int n = blockIdx.y;
    const int cpg = C / groups;
    // fixed-order reduction of the split statistics: 8 strided partial sums per group, then those 8 in order
    __shared__ float s_pa[64][8], s_pb[64][8];
    for (int t = threadIdx.x; t < groups * 8; t += blockDim.x) {
        const int g = t >> 3, k = t & 7;
        float a = 0.f, b = 0.f;
        for (int s = k; s < splits; s += 8) {
            const float2 in = *reinterpret_cast<const float2*>(partial + (((size_t)n * splits + s) * groups + g) * 2);
            a += in.x, b += in.y;
        }
        s_pa[g][k] = a, s_pb[g][k] = b;
    }
    __syncthreads();
    for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) a += s_pa[g][k], b += s_pb[g][k];
        const float inv = 1.0f / ((float)cpg * stat_hw);  // stat_hw = pixels the sums cover (hw, or the whole image's)
        const float mean = a * inv;
        const float var = fmaxf(b * inv - mean * mean, 0.f);
        s_mean[g] = mean, s_rstd[g] = rsqrtf(var + eps);
    }
    __syncthreads();
    float sc[VEC], sh[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const int g = (ch + j) / cpg;
        const float w = gamma[ch + j] * s_rstd[g];
        sc[j] = w, sh[j] = beta[ch + j] - s_mean[g] * w;
    }
    const bool second = ch >= c1;
    const void* src = second ? x2 : x1;
    const int cs = second ? c2 : c1;
    const size_t off = (size_t)n * hw * cs + (second ? ch - c1 : ch);
    const int per = (hw + chunks - 1) / chunks;
    const int p0 = blockIdx.x * per;
    const int p1 = min(hw, p0 + per);
    const bool f16 = out_f16 != 0;
    auto emit = [&](int p, const float (&v)[VEC]) {
        const size_t o = ((size_t)n * hw + p) * C + ch;
        if (raw) storev_h<VEC>(raw, o, v, f16);
        float r[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            r[j] = v[j] * sc[j] + sh[j];
            if (silu) r[j] = silu_f(r[j]);
        }
        storev_h<VEC>(y, o, r, f16);
    };
#ifdef B200SD_GN_APPLY_U
    constexpr int U = B200SD_GN_APPLY_U;
#else
    constexpr int U = 4;  // loads in flight per thread (16 bytes each with VEC = 8): 2 left the SM short of bytes in flight
#endif
    int p = p0 + pl;
    for (; p + (U - 1) * lanes < p1; p += U * lanes) {
        float v[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u) loadv<VEC>(src, off + (size_t)(p + u * lanes) * cs, dt, v[u]);
#pragma unroll
        for (int u = 0; u < U; ++u) emit(p + u * lanes, v[u]);
    }
    for (; p < p1; p += lanes) {
        float v[VEC];
        loadv<VEC>(src, off + (size_t)p * cs, dt, v);
        emit(p, v);
    }
}

static int gn_splits(int n_img, int hw) {
    // ~8 CTAs per SM overall: one resident wave of small CTAs keeps >100 KB of loads in flight per SM and leaves no
    // 3-vs-2 CTA imbalance between SMs (measured at 2 per SM: 2.6 TB/s)
    int want = (8 * num_sms() + n_img - 1) / n_img;
    int max_by_pixels = (hw + 31) / 32;              // at least 32 pixels per split
    int s = want < max_by_pixels ? want : max_by_pixels;
    if (s < 1) s = 1;
    if (s > 256) s = 256;
    return s;
}

// Per-(image, group) sums from the split statistics: out[n][g] = sum over splits, in split order.
__global__ void gn_reduce_splits_kernel(const float* __restrict__ partial, int splits, int groups, float* __restrict__ out) {
    pdl_sync();
    const int n = blockIdx.x;
    for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        float a = 0.f, b = 0.f;
        for (int s = 0; s < splits; ++s) {
            const float2 in = *reinterpret_cast<const float2*>(partial + (((size_t)n * splits + s) * groups + g) * 2);
            a += in.x, b += in.y;
        }
        out[((size_t)n * groups + g) * 2] = a, out[((size_t)n * groups + g) * 2 + 1] = b;
    }
}

// Per-(image, group) sums from the column sums a producing GEMM's epilogue left behind (gemm.cu, staged_store):
// t[n][part][c] = (sum, sum of squares) of column c over one 32-row slice of a tile.  One CTA per (group, image);
// every thread adds a fixed subset of its group's [parts x channels-per-group] block in a fixed order, then the 128
// per-thread sums are added by a fixed tree: the result does not depend on scheduling.
//
// gn_tile_block_sum adds the [parts x w] block of (sum, sum of squares) pairs at `base` (row stride `ld` pairs) into (a, b): thread i of 128
// walks a fixed set of elements in a fixed order.  Even widths go two channels (one 16-byte load) at a time, four rows in
// flight per thread: the table was written a moment ago by the producing GEMM and sits in L2, so the loop is bound by
// load latency, not bandwidth.
__device__ __forceinline__ void gn_tile_block_sum(const float2* __restrict__ base, int parts, int w, int ld, float& a,
                                                  float& b) {
    if ((w & 1) == 0 && (ld & 1) == 0 && (reinterpret_cast<uintptr_t>(base) & 15) == 0) {
        const int wp = w >> 1;
        const int rows = 128 / wp;  // rows covered per pass (w <= 256)
        const int r0 = threadIdx.x / wp, cp = threadIdx.x - r0 * wp;
        if (rows == 0 || r0 >= rows) {
            if (rows != 0) return;
        } else {
            const float4* src = reinterpret_cast<const float4*>(base + 2 * cp);
            const size_t ld4 = (size_t)ld >> 1;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
            int part = r0;
            for (; part + 3 * rows < parts; part += 4 * rows) {
                const float4 v0 = src[(size_t)part * ld4], v1 = src[(size_t)(part + rows) * ld4];
                const float4 v2 = src[(size_t)(part + 2 * rows) * ld4], v3 = src[(size_t)(part + 3 * rows) * ld4];
                a0 += v0.x + v0.z, b0 += v0.y + v0.w, a1 += v1.x + v1.z, b1 += v1.y + v1.w;
                a2 += v2.x + v2.z, b2 += v2.y + v2.w, a3 += v3.x + v3.z, b3 += v3.y + v3.w;
            }
            for (; part < parts; part += rows) {
                const float4 v0 = src[(size_t)part * ld4];
                a0 += v0.x + v0.z, b0 += v0.y + v0.w;
            }
            a += (a0 + a1) + (a2 + a3), b += (b0 + b1) + (b2 + b3);
            return;
        }
    }
    for (int i = threadIdx.x; i < parts * w; i += 128) {
        const int part = i / w, c = i - part * w;
        const float2 v = base[(size_t)part * ld + c];
        a += v.x, b += v.y;
    }
}

__global__ void __launch_bounds__(128)
    gn_tile_sums_kernel(const float2* __restrict__ t1, int parts1, int c1, const float2* __restrict__ t2, int parts2,
                        int c2, int groups, float* __restrict__ out) {
    pdl_sync();
    __shared__ float sa[128], sb[128];
    const int g = blockIdx.x, n = blockIdx.y;
    const int cpg = (c1 + c2) / groups;
    const int ch0 = g * cpg, ch1 = ch0 + cpg;
    float a = 0.f, b = 0.f;
    // channels [ch0, ch1) of the concatenation: [ch0, min(ch1, c1)) from the first producer, the rest from the second
    const int e1 = min(ch1, c1);
    if (ch0 < e1) gn_tile_block_sum(t1 + (size_t)n * parts1 * c1 + ch0, parts1, e1 - ch0, c1, a, b);
    const int b2 = max(ch0, c1);
    if (b2 < ch1) gn_tile_block_sum(t2 + (size_t)n * parts2 * c2 + (b2 - c1), parts2, ch1 - b2, c2, a, b);
    sa[threadIdx.x] = a, sb[threadIdx.x] = b;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (threadIdx.x < o) sa[threadIdx.x] += sa[threadIdx.x + o], sb[threadIdx.x] += sb[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[((size_t)n * groups + g) * 2] = sa[0], out[((size_t)n * groups + g) * 2 + 1] = sb[0];
}

// =============================================================================================== LayerNorm
// one warp per row, whole row held in registers (cols <= 1536): exact two-pass mean / variance.
// LW = lanes per row: 32 (a warp per row) or 16 (two rows per warp, for rows of at most 16 * 4 * MAXV columns: the 320
// columns of the 96x96 level are 80 four-element vectors = 5 per lane of a half warp, where a full warp would idle half of
// its lanes on the third vector)
template <int MAXV, int DT, bool OF16, int LW>
__global__ void layernorm_kernel(const void* __restrict__ x, long long rows, int cols, float eps,
                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                 void* __restrict__ y) {
    constexpr int dt = DT;          // formats fixed at compile time (see gn_apply_kernel): the generic version spent 81 % of
    constexpr int out_f16 = OF16;   // its issue slots, most of them on decoding both 16-bit formats per element
    constexpr int RPW = 32 / LW;    // rows per warp
    pdl_sync();
    const int lane = threadIdx.x & (LW - 1);
    const long long row = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + ((threadIdx.x & 31) / LW);
    const bool live = row < rows;   // (the other half of a warp may still hold a row: no early return before the shuffles)
    const int nvec = live ? cols >> 2 : 0;
    const size_t base = (size_t)row * cols;
    float4 v[MAXV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int idx = lane + i * LW;
        if (idx < nvec) {
            v[i] = load4(x, base + idx * 4, dt);
            s += v[i].x + v[i].y + v[i].z + v[i].w;
        }
    }
#pragma unroll
    for (int o = LW / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / (float)cols;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int idx = lane + i * LW;
        if (idx < nvec) {
            float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            q += a * a + b * b + c * c + d * d;
        }
    }
#pragma unroll
    for (int o = LW / 2; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / (float)cols + eps);
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int idx = lane + i * LW;
        if (idx < nvec) {
            float4 g = *reinterpret_cast<const float4*>(gamma + idx * 4);
            float4 b = *reinterpret_cast<const float4*>(beta + idx * 4);
            float4 r = make_float4((v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y,
                                   (v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w);
            store4_h(y, base + idx * 4, r, out_f16 != 0);
        }
    }
}

// =============================================================================================== token embeddings
// out[r, :] = token_table[ids[r], :] + position_table[r % seq_len, :]; one thread per 4 channels
__global__ void embed_tokens_kernel(const int* __restrict__ ids, const float* __restrict__ tok,
                                    const float* __restrict__ pos, float* __restrict__ out, long long rows, int seq_len,
                                    int dim) {
    pdl_sync();
    const int vpr = dim >> 2;
    const long long total = rows * vpr;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / vpr;
        const int v = (int)(i - r * vpr);
        const float4 a = *reinterpret_cast<const float4*>(tok + (size_t)ids[r] * dim + 4 * v);
        const float4 b = *reinterpret_cast<const float4*>(pos + (size_t)(r % seq_len) * dim + 4 * v);
        *reinterpret_cast<float4*>(out + (size_t)r * dim + 4 * v) = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    }
}

// =============================================================================================== im2col 3x3
// one thread per 8 consecutive k of one output pixel
__global__ void im2col3x3_kernel(const void* __restrict__ x, int dt, int n_img, int n_src, int in_h, int in_w,
                                 int cin, int stride, int out_h, int out_w, int k_pad, void* __restrict__ out,
                                 int out_f16, int pad_lo) {
    pdl_sync();
    const int kchunks = k_pad >> 3;
    const long long total = (long long)n_img * out_h * out_w * kchunks;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int kc = (int)(i % kchunks);
        const long long m = i / kchunks;
        const int ox = (int)(m % out_w);
        const int oy = (int)((m / out_w) % out_h);
        const int n = (int)(m / ((long long)out_w * out_h));
        const size_t src = (size_t)(n % n_src) * in_h * in_w * cin;
        float v[8];
        const int k0 = kc << 3;
        if ((cin & 7) == 0) {
            const int tap = k0 / cin, c = k0 - tap * cin;
            const int iy = oy * stride + tap / 3 - pad_lo, ix = ox * stride + tap % 3 - pad_lo;
            if (tap < 9 && iy >= 0 && iy < in_h && ix >= 0 && ix < in_w) {
                const size_t p = src + ((size_t)iy * in_w + ix) * cin + c;
                const float4 a = load4(x, p, dt), b = load4(x, p + 4, dt);
                v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = 0.f;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int k = k0 + j;
                const int tap = k / cin, c = k - tap * cin;
                const int iy = oy * stride + tap / 3 - pad_lo, ix = ox * stride + tap % 3 - pad_lo;
                float t = 0.f;
                if (tap < 9 && iy >= 0 && iy < in_h && ix >= 0 && ix < in_w)
                    t = load1(x, src + ((size_t)iy * in_w + ix) * cin + c, dt);
                v[j] = t;
            }
        }
        store8_h(out, (size_t)m * k_pad + k0, v, out_f16 != 0);
    }
}

// =============================================================================================== upsample 2x
__global__ void upsample2x_kernel(const void* __restrict__ x, int dt, int n_img, int h, int w, int c,
                                  void* __restrict__ out, int out_f16) {
    pdl_sync();
    const int cv = c >> 3;
    const long long total = (long long)n_img * (2 * h) * (2 * w) * cv;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int v8 = (int)(i % cv);
        const long long pix = i / cv;
        const int ox = (int)(pix % (2 * w));
        const int oy = (int)((pix / (2 * w)) % (2 * h));
        const int n = (int)(pix / ((long long)4 * w * h));
        const size_t p = (((size_t)n * h + (oy >> 1)) * w + (ox >> 1)) * c + v8 * 8;
        const float4 a = load4(x, p, dt), b = load4(x, p + 4, dt);
        const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        store8_h(out, (size_t)pix * c + v8 * 8, v, out_f16 != 0);
    }
}

// =============================================================================================== K4
__global__ void cfg_sched_kernel(const void* __restrict__ eps2, int dt, float* __restrict__ x,
                                 float* __restrict__ x0_prev, const float* __restrict__ coef,
                                 const int* __restrict__ step_index, long long n) {
    pdl_sync();
    const int step = step_index ? *step_index : 0;
    const float g = coef[step * 6 + 0], cx = coef[step * 6 + 1], cm = coef[step * 6 + 2], cp = coef[step * 6 + 3],
                ax = coef[step * 6 + 4], am = coef[step * 6 + 5];
    const long long nv = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long long)gridDim.x * blockDim.x) {
        const float4 u = load4(eps2, (size_t)i * 4, dt);
        const float4 c = load4(eps2, (size_t)(n + i * 4), dt);
        const float4 xv = *reinterpret_cast<const float4*>(x + i * 4);
        const float4 m = make_float4(u.x + g * (c.x - u.x), u.y + g * (c.y - u.y), u.z + g * (c.z - u.z),
                                     u.w + g * (c.w - u.w));
        float4 o = make_float4(cx * xv.x + cm * m.x, cx * xv.y + cm * m.y, cx * xv.z + cm * m.z,
                               cx * xv.w + cm * m.w);
        if (x0_prev) {
            const float4 pv = *reinterpret_cast<const float4*>(x0_prev + i * 4);
            o.x += cp * pv.x, o.y += cp * pv.y, o.z += cp * pv.z, o.w += cp * pv.w;
            *reinterpret_cast<float4*>(x0_prev + i * 4) =
                make_float4(ax * xv.x + am * m.x, ax * xv.y + am * m.y, ax * xv.z + am * m.z, ax * xv.w + am * m.w);
        }
        *reinterpret_cast<float4*>(x + i * 4) = o;
    }
}

// =============================================================================================== K5
// one warp per output column, rows processed 8 at a time from shared memory
__global__ void small_linear_kernel(const float* __restrict__ x, int rows, int k, int n, int sinusoid_dim,
                                    int silu_in, int silu_out, const uint16_t* __restrict__ w, int w_f16,
                                    const float* __restrict__ b, float* __restrict__ y) {
    pdl_sync();
    extern __shared__ float xs[];  // [8][k]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * (blockDim.x >> 5) + warp;
    const bool f16 = w_f16 != 0;
    for (int r0 = 0; r0 < rows; r0 += 8) {
        const int nr = min(8, rows - r0);
        __syncthreads();
        for (int i = threadIdx.x; i < nr * k; i += blockDim.x) {
            const int r = i / k, kk = i - r * k;
            float v;
            if (sinusoid_dim > 0) {
                const int half = sinusoid_dim >> 1;
                const int f = kk < half ? kk : kk - half;
                const float freq = expf(-9.210340371976184f * (float)f / (float)half);  // ln(10000)
                const float a = x[r0 + r] * freq;
                v = kk < half ? cosf(a) : sinf(a);
            } else {
                v = x[(size_t)(r0 + r) * k + kk];
            }
            xs[i] = silu_in ? silu_f(v) : v;
        }
        __syncthreads();
        if (col < n) {
            float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const uint16_t* wr = w + (size_t)col * k;
            for (int kk = lane * 8; kk < k; kk += 256) {
                const uint4 wv = *reinterpret_cast<const uint4*>(wr + kk);
                const float wf[8] = {h_lo(wv.x, f16), h_hi(wv.x, f16), h_lo(wv.y, f16), h_hi(wv.y, f16),
                                     h_lo(wv.z, f16), h_hi(wv.z, f16), h_lo(wv.w, f16), h_hi(wv.w, f16)};
                for (int r = 0; r < nr; ++r) {
                    const float* xr = xs + r * k + kk;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[r] += wf[j] * xr[j];
                }
            }
            for (int r = 0; r < nr; ++r) {
                float a = acc[r];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
                if (lane == 0) {
                    a += b ? b[col] : 0.f;
                    y[(size_t)(r0 + r) * n + col] = silu_out ? silu_f(a) : a;
                }
            }
        }
    }
}

// =============================================================================================== row softmax
__global__ void softmax_rows_kernel(const void* __restrict__ x, int dt, long long ld_in, void* __restrict__ y,
                                    int out_f16, long long ld_out, int cols, float scale) {
    pdl_sync();
    extern __shared__ float row[];
    __shared__ float red[32];
    const size_t in0 = (size_t)blockIdx.x * ld_in, out0 = (size_t)blockIdx.x * ld_out;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    float mx = -INFINITY;
    for (int i = tid * 4; i < cols; i += blockDim.x * 4) {
        float4 v = load4(x, in0 + i, dt);
        v.x *= scale, v.y *= scale, v.z *= scale, v.w *= scale;
        *reinterpret_cast<float4*>(row + i) = v;
        mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int i = 1; i < nw; ++i) mx = fmaxf(mx, red[i]);
    __syncthreads();
    float sum = 0.f;
    for (int i = tid; i < cols; i += blockDim.x) {
        const float e = __expf(row[i] - mx);
        row[i] = e;
        sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int i = 0; i < nw; ++i) sum += red[i];
    const float inv = 1.0f / sum;
    for (int i = tid * 4; i < cols; i += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4*>(row + i);
        store4_h(y, out0 + i, make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv), out_f16 != 0);
    }
}

// =============================================================================================== K7 / cast
__global__ void denorm_u8_kernel(const void* __restrict__ x, int dt, uint8_t* __restrict__ out, long long n) {
    pdl_sync();
    const long long nv = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long long)gridDim.x * blockDim.x) {
        const float4 v = load4(x, (size_t)i * 4, dt);
        // ((x / 2 + 0.5).clamp(0, 1) * 255).round(): one rounding per operation as in the reference's fp32 tensor ops
        // (explicit _rn intrinsics keep the compiler from contracting anything), rintf = round half to even = torch.round
        auto q = [](float x) {
            const float t = fminf(fmaxf(__fadd_rn(__fmul_rn(x, 0.5f), 0.5f), 0.f), 1.f);
            return (unsigned char)rintf(__fmul_rn(t, 255.f));
        };
        uchar4 o;
        o.x = q(v.x), o.y = q(v.y), o.z = q(v.z), o.w = q(v.w);
        *reinterpret_cast<uchar4*>(out + i * 4) = o;
    }
}

__global__ void cast_kernel(const void* __restrict__ x, int dt_in, void* __restrict__ y, int dt_out, long long n) {
    pdl_sync();
    const long long nv = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (long long)gridDim.x * blockDim.x) {
        const float4 v = load4(x, (size_t)i * 4, dt_in);
        if (dt_out == B200SD_F32)
            *reinterpret_cast<float4*>(static_cast<float*>(y) + i * 4) = v;
        else
            store4_h(y, (size_t)i * 4, v, dt_out == B200SD_F16);
    }
}

static int ew_grid(long long work_items, int block) {
    long long blocks = (work_items + block - 1) / block;
    long long cap = (long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace b200sd

using namespace b200sd;

#define ALIGNED16(p) ((reinterpret_cast<uintptr_t>(p) & 15) == 0)
#define ANY_DTYPE(d) ((d) == B200SD_BF16 || (d) == B200SD_F32 || (d) == B200SD_F16)
#define H_DTYPE(d) ((d) == B200SD_BF16 || (d) == B200SD_F16)

extern "C" int64_t b200sd_groupnorm_workspace_bytes(const b200sd_groupnorm_params* p) {
    if (!p || p->n_img <= 0 || p->hw <= 0 || p->groups <= 0) return 0;
    return (int64_t)p->n_img * gn_splits(p->n_img, p->hw) * p->groups * 2 * sizeof(float);
}

namespace b200sd {

// argument checks shared by the two GroupNorm entry points
static int32_t gn_check(const b200sd_groupnorm_params* p, const char* who) {
    B200SD_CHECK_ARG(p && p->x1 && (p->sums_out || (p->y && p->gamma && p->beta)), "%s: null pointer", who);
    const int c2 = p->x2 ? p->c2 : 0;
    const int C = p->c1 + c2;
    B200SD_CHECK_ARG(p->n_img > 0 && p->hw > 0, "%s: bad shape", who);
    B200SD_CHECK_ARG(p->groups > 0 && p->groups <= 64 && C % p->groups == 0, "%s: C=%d groups=%d", who, C, p->groups);
    B200SD_CHECK_ARG(p->c1 % 4 == 0 && c2 % 4 == 0 && C / 4 <= 1024, "%s: channels must be multiples of 4, C<=4096", who);
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype) && H_DTYPE(p->out_dtype), "%s: bad in_dtype / out_dtype", who);
    B200SD_CHECK_ARG(ALIGNED16(p->x1) && (!p->x2 || ALIGNED16(p->x2)) && ALIGNED16(p->y) && (!p->raw || ALIGNED16(p->raw)),
                     "%s: pointers must be 16-byte aligned", who);
    return B200SD_OK;
}

// The launches of GroupNorm: split statistics into `scratch` (n_img * splits * groups * 2 floats), then reduce +
// normalise.  Both kernels accept any split count >= 1.  `stats` / `apply` select the halves: statistics only
// (p->sums_out receives the per-(image, group) sums), apply only (p->sums given: they ARE the one split), or both.
static int32_t gn_launch(const b200sd_groupnorm_params* p, int splits, float* scratch, cudaStream_t s, const char* who,
                         bool stats = true, bool apply = true) {
    const int c2 = p->x2 ? p->c2 : 0;
    const int C = p->c1 + c2;
    const bool vec8 = p->c1 % 8 == 0 && c2 % 8 == 0;  // 8 channels per thread: 16-byte accesses on the 16-bit side
    const int vpp = C / (vec8 ? 8 : 4);
    int lanes = 256 / vpp;
    if (lanes < 1) lanes = 1;
    const int threads = vpp * lanes;
    const size_t smem = (size_t)2 * lanes * C * sizeof(float);
    const int chunks = gn_splits(p->n_img, p->hw) * 2;
    dim3 g1(splits, p->n_img), g2(chunks, p->n_img);
    int launches = 0;
    const int dt = p->in_dtype, f16 = p->out_dtype == B200SD_F16;
    if (stats) {
        if (vec8) {
            B200SD_LAUNCH("gn_stats_kernel", (gn_stats_kernel<8>), g1, dim3(threads), smem, s, p->x1, p->c1, p->x2, c2, dt, p->hw,
                          p->groups, splits, scratch);
        } else {
            B200SD_LAUNCH("gn_stats_kernel", (gn_stats_kernel<4>), g1, dim3(threads), smem, s, p->x1, p->c1, p->x2, c2, dt, p->hw,
                          p->groups, splits, scratch);
        }
        ++launches;
    }
    if (stats && !apply) {
        B200SD_LAUNCH("gn_reduce_splits_kernel", gn_reduce_splits_kernel, dim3(p->n_img), dim3(64), 0, s, scratch, splits,
                      p->groups, p->sums_out);
        ++launches;
    }
    if (apply) {
        const float* part = stats ? scratch : p->sums;
        const int nsplit = stats ? splits : 1;
        const float stat_hw = (float)((!stats && p->stat_hw > 0) ? p->stat_hw : (int64_t)p->hw);
#define B200SD_GN_APPLY(V, D, O, S)                                                                                       \
    B200SD_LAUNCH("gn_apply_kernel", (gn_apply_kernel<V, D, O, S>), g2, dim3(threads), 0, s, p->x1, p->c1, p->x2, c2, p->hw, \
                  p->groups, nsplit, part, p->eps, p->gamma, p->beta, p->y, p->raw, chunks, stat_hw)
#define B200SD_GN_APPLY_S(V, D, O)                                      \
    do {                                                                \
        if (p->silu) B200SD_GN_APPLY(V, D, O, true);                    \
        else B200SD_GN_APPLY(V, D, O, false);                           \
    } while (0)
#define B200SD_GN_APPLY_O(V, D)                                         \
    do {                                                                \
        if (f16) B200SD_GN_APPLY_S(V, D, true);                         \
        else B200SD_GN_APPLY_S(V, D, false);                            \
    } while (0)
#define B200SD_GN_APPLY_D(V)                                            \
    do {                                                                \
        if (dt == B200SD_F32) B200SD_GN_APPLY_O(V, B200SD_F32);         \
        else if (dt == B200SD_F16) B200SD_GN_APPLY_O(V, B200SD_F16);    \
        else B200SD_GN_APPLY_O(V, B200SD_BF16);                         \
    } while (0)
        if (vec8) B200SD_GN_APPLY_D(8);
        else B200SD_GN_APPLY_D(4);
#undef B200SD_GN_APPLY_D
#undef B200SD_GN_APPLY_O
#undef B200SD_GN_APPLY_S
#undef B200SD_GN_APPLY
        ++launches;
    }
    count_launch(launches);
    B200SD_CHECK_LAUNCH(who);
    return B200SD_OK;
}

}  // namespace b200sd

extern "C" int32_t b200sd_groupnorm(const b200sd_groupnorm_params* p, b200sd_stream_t stream) {
    const int32_t rc = gn_check(p, "groupnorm");
    if (rc != B200SD_OK) return rc;
    B200SD_CHECK_ARG(!(p->sums && p->sums_out), "groupnorm: sums and sums_out are exclusive");
    if (p->sums) {  // statistics supplied by the caller (a producer's epilogue, or an all-reduce over GPUs): apply only
        B200SD_CHECK_ARG(p->stat_hw == 0 || p->stat_hw >= p->hw, "groupnorm: stat_hw %lld < hw %d", (long long)p->stat_hw, p->hw);
        return gn_launch(p, 1, nullptr, as_stream(stream), "groupnorm (apply)", false, true);
    }
    B200SD_CHECK_ARG(p->workspace, "groupnorm: null workspace");
    const int splits = gn_splits(p->n_img, p->hw);
    const int64_t need = (int64_t)p->n_img * splits * p->groups * 2 * sizeof(float);
    if (p->workspace_bytes < need)
        return set_error(B200SD_ERR_WORKSPACE, "groupnorm: workspace %lld < %lld bytes", (long long)p->workspace_bytes,
                         (long long)need);
    if (p->sums_out) return gn_launch(p, splits, p->workspace, as_stream(stream), "groupnorm (statistics)", true, false);
    return gn_launch(p, splits, p->workspace, as_stream(stream), "groupnorm");
}

extern "C" int32_t b200sd_gn_tile_sums(const b200sd_gn_tile_sums_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->tile_sums1 && p->sums, "gn_tile_sums: null pointer");
    const int c2 = p->tile_sums2 ? p->c2 : 0;
    B200SD_CHECK_ARG(p->n_img > 0 && p->parts1 > 0 && p->c1 > 0 && (!p->tile_sums2 || (p->parts2 > 0 && p->c2 > 0)),
                     "gn_tile_sums: bad shape");
    B200SD_CHECK_ARG(p->groups > 0 && p->groups <= 64 && (p->c1 + c2) % p->groups == 0, "gn_tile_sums: C=%d groups=%d",
                     p->c1 + c2, p->groups);
    B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(p->tile_sums1) & 7) == 0 &&
                         (reinterpret_cast<uintptr_t>(p->tile_sums2) & 7) == 0, "gn_tile_sums: tables must be 8-byte aligned");
    B200SD_LAUNCH("gn_tile_sums_kernel", gn_tile_sums_kernel, dim3(p->groups, p->n_img), dim3(128), 0, as_stream(stream), 
        reinterpret_cast<const float2*>(p->tile_sums1), p->parts1, p->c1, reinterpret_cast<const float2*>(p->tile_sums2),
        p->parts2, c2, p->groups, p->sums);
    count_launch();
    B200SD_CHECK_LAUNCH("gn_tile_sums");
    return B200SD_OK;
}

extern "C" int32_t b200sd_embed_tokens(const b200sd_embed_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->ids && p->token_table && p->position_table && p->out, "embed_tokens: null pointer");
    B200SD_CHECK_ARG(p->dim > 0 && p->dim % 4 == 0 && p->seq_len > 0 && p->vocab > 0 && p->rows >= 0,
                     "embed_tokens: bad shape (dim %d must be a positive multiple of 4)", p->dim);
    B200SD_CHECK_ARG(ALIGNED16(p->token_table) && ALIGNED16(p->position_table) && ALIGNED16(p->out),
                     "embed_tokens: tables and output must be 16-byte aligned");
    if (p->rows == 0) return B200SD_OK;
    B200SD_LAUNCH("embed_tokens_kernel", embed_tokens_kernel, dim3(ew_grid(p->rows * (p->dim / 4), 256)), dim3(256), 0, as_stream(stream), 
        p->ids, p->token_table, p->position_table, p->out, p->rows, p->seq_len, p->dim);
    count_launch();
    B200SD_CHECK_LAUNCH("embed_tokens");
    return B200SD_OK;
}

extern "C" int32_t b200sd_layernorm(const b200sd_layernorm_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->y && p->gamma && p->beta, "layernorm: null pointer");
    B200SD_CHECK_ARG(p->cols % 4 == 0 && p->cols <= 1536 && p->cols > 0, "layernorm: cols=%d (need %%4==0, <=1536)",
                     p->cols);
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype) && H_DTYPE(p->out_dtype), "layernorm: bad in_dtype / out_dtype");
    B200SD_CHECK_ARG(ALIGNED16(p->x) && ALIGNED16(p->y) && ALIGNED16(p->gamma) && ALIGNED16(p->beta),
                     "layernorm: pointers must be 16-byte aligned");
    if (p->rows <= 0) return B200SD_OK;
    const int wpb = 8;
    // registers per thread follow the row length (4 * MAXV floats), so short rows run at full occupancy
    const int nvec4 = p->cols / 4;
    const bool half_warp = nvec4 <= 80;           // two rows per warp, 16 lanes each, up to 5 vectors per lane
    const int nv = (nvec4 + 31) / 32;
    const long long rows_per_block = (long long)wpb * (half_warp ? 2 : 1);
    const unsigned grid = (unsigned)((p->rows + rows_per_block - 1) / rows_per_block);
#define B200SD_LN_LAUNCH_T(MV, D, O, LW)                                                                             \
    B200SD_LAUNCH("layernorm_kernel", (layernorm_kernel<MV, D, O, LW>), dim3(grid), dim3(wpb * 32), 0, as_stream(stream), p->x, p->rows, \
                  p->cols, p->eps, p->gamma, p->beta, p->y)
#define B200SD_LN_LAUNCH_O(MV, D, LW)                                                 \
    do {                                                                              \
        if (p->out_dtype == B200SD_F16) B200SD_LN_LAUNCH_T(MV, D, true, LW);          \
        else B200SD_LN_LAUNCH_T(MV, D, false, LW);                                    \
    } while (0)
#define B200SD_LN_LAUNCH(MV, LW)                                                      \
    do {                                                                              \
        if (p->in_dtype == B200SD_F32) B200SD_LN_LAUNCH_O(MV, B200SD_F32, LW);        \
        else if (p->in_dtype == B200SD_F16) B200SD_LN_LAUNCH_O(MV, B200SD_F16, LW);   \
        else B200SD_LN_LAUNCH_O(MV, B200SD_BF16, LW);                                 \
    } while (0)
    if (half_warp) B200SD_LN_LAUNCH(5, 16);
    else if (nv <= 5) B200SD_LN_LAUNCH(5, 32);
    else if (nv <= 10) B200SD_LN_LAUNCH(10, 32);
    else B200SD_LN_LAUNCH(12, 32);
#undef B200SD_LN_LAUNCH_T
#undef B200SD_LN_LAUNCH_O
#undef B200SD_LN_LAUNCH
    count_launch();
    B200SD_CHECK_LAUNCH("layernorm");
    return B200SD_OK;
}

extern "C" int32_t b200sd_im2col3x3(const b200sd_im2col_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->out, "im2col: null pointer");
    B200SD_CHECK_ARG(p->stride == 1 || p->stride == 2, "im2col: stride must be 1 or 2");
    B200SD_CHECK_ARG(p->k_pad % 8 == 0 && p->k_pad >= 9 * p->cin, "im2col: k_pad=%d cin=%d", p->k_pad, p->cin);
    B200SD_CHECK_ARG(p->n_src > 0 && p->n_img > 0, "im2col: bad batch");
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype) && H_DTYPE(p->out_dtype), "im2col: bad in_dtype / out_dtype");
    B200SD_CHECK_ARG(ALIGNED16(p->x) && ALIGNED16(p->out), "im2col: pointers must be 16-byte aligned");
    B200SD_CHECK_ARG(p->pad_lo == 0 || p->pad_lo == 1, "im2col: pad_lo must be 0 or 1");
    B200SD_CHECK_ARG(p->in_h + p->pad_lo >= 2 && p->in_w + p->pad_lo >= 2, "im2col: image smaller than the filter");
    const int out_h = (p->in_h + p->pad_lo + 1 - 3) / p->stride + 1, out_w = (p->in_w + p->pad_lo + 1 - 3) / p->stride + 1;
    const long long total = (long long)p->n_img * out_h * out_w * (p->k_pad / 8);
    B200SD_LAUNCH("im2col3x3_kernel", im2col3x3_kernel, dim3(ew_grid(total, 256)), dim3(256), 0, as_stream(stream), 
        p->x, p->in_dtype, p->n_img, p->n_src, p->in_h, p->in_w, p->cin, p->stride, out_h, out_w, p->k_pad, p->out,
        p->out_dtype == B200SD_F16, p->pad_lo);
    count_launch();
    B200SD_CHECK_LAUNCH("im2col3x3");
    return B200SD_OK;
}

extern "C" int32_t b200sd_upsample2x(const b200sd_upsample_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->out, "upsample2x: null pointer");
    B200SD_CHECK_ARG(p->c % 8 == 0, "upsample2x: c must be a multiple of 8");
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype) && H_DTYPE(p->out_dtype), "upsample2x: bad in_dtype / out_dtype");
    B200SD_CHECK_ARG(ALIGNED16(p->x) && ALIGNED16(p->out), "upsample2x: pointers must be 16-byte aligned");
    const long long total = (long long)p->n_img * 4 * p->h * p->w * (p->c / 8);
    B200SD_LAUNCH("upsample2x_kernel", upsample2x_kernel, dim3(ew_grid(total, 256)), dim3(256), 0, as_stream(stream), p->x, p->in_dtype, p->n_img, p->h, p->w,
                                                                         p->c, p->out, p->out_dtype == B200SD_F16);
    count_launch();
    B200SD_CHECK_LAUNCH("upsample2x");
    return B200SD_OK;
}

extern "C" int32_t b200sd_cfg_sched_step(const b200sd_cfg_sched_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->eps2 && p->x && p->coef, "cfg_sched_step: null pointer");
    B200SD_CHECK_ARG(p->n > 0 && p->n % 4 == 0, "cfg_sched_step: n must be a positive multiple of 4");
    B200SD_CHECK_ARG(ANY_DTYPE(p->eps_dtype), "cfg_sched_step: bad eps_dtype");
    B200SD_CHECK_ARG(ALIGNED16(p->eps2) && ALIGNED16(p->x) && (!p->x0_prev || ALIGNED16(p->x0_prev)) &&
                         (p->eps_dtype == B200SD_F32 || (p->n % 8 == 0)),
                     "cfg_sched_step: pointers must be 16-byte aligned (16-bit eps: n %% 8 == 0)");
    B200SD_LAUNCH("cfg_sched_kernel", cfg_sched_kernel, dim3(ew_grid(p->n / 4, 256)), dim3(256), 0, as_stream(stream), p->eps2, p->eps_dtype, p->x, p->x0_prev,
                                                                           p->coef, p->step_index, p->n);
    count_launch();
    B200SD_CHECK_LAUNCH("cfg_sched_step");
    return B200SD_OK;
}

extern "C" int32_t b200sd_small_linear(const b200sd_small_linear_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->w && p->y, "small_linear: null pointer");
    B200SD_CHECK_ARG(p->rows > 0 && p->rows <= 64 && p->k % 8 == 0 && p->k > 0 && p->k <= 4096 && p->n > 0,
                     "small_linear: rows=%d k=%d n=%d", p->rows, p->k, p->n);
    B200SD_CHECK_ARG(p->sinusoid_dim == 0 || p->sinusoid_dim == p->k, "small_linear: sinusoid_dim must be 0 or k");
    B200SD_CHECK_ARG(H_DTYPE(p->w_dtype) && ALIGNED16(p->w), "small_linear: w must be bf16/fp16, 16-byte aligned");
    const int wpb = 8;
    const int grid = (p->n + wpb - 1) / wpb;
    const size_t smem = (size_t)8 * p->k * sizeof(float);
    static bool attr_set[64] = {false};  // the attribute is per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > 48 * 1024 && dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaFuncSetAttribute(small_linear_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 4096 * 4);
        attr_set[dev] = true;
    }
    B200SD_LAUNCH("small_linear_kernel", small_linear_kernel, dim3(grid), dim3(wpb * 32), smem, as_stream(stream), 
        p->x, p->rows, p->k, p->n, p->sinusoid_dim, p->silu_in, p->silu_out, (const uint16_t*)p->w,
        p->w_dtype == B200SD_F16, p->b, p->y);
    count_launch();
    B200SD_CHECK_LAUNCH("small_linear");
    return B200SD_OK;
}

extern "C" int32_t b200sd_softmax_rows(const b200sd_softmax_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->y, "softmax_rows: null pointer");
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype) && H_DTYPE(p->out_dtype), "softmax_rows: bad in_dtype / out_dtype");
    B200SD_CHECK_ARG(p->cols > 0 && p->cols % 4 == 0 && p->ld_in % 4 == 0 && p->ld_out % 4 == 0 && p->cols <= 49152 &&
                         ALIGNED16(p->x) && (reinterpret_cast<uintptr_t>(p->y) & 7) == 0,
                     "softmax_rows: cols=%d ld_in=%lld ld_out=%lld (need %%4==0, cols<=49152, aligned)", p->cols,
                     (long long)p->ld_in, (long long)p->ld_out);
    if (p->rows <= 0) return B200SD_OK;
    const size_t smem = (size_t)p->cols * sizeof(float);
    static bool attr_set[64] = {false};  // the attribute is per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaFuncSetAttribute(softmax_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152 * 4);
        attr_set[dev] = true;
    }
    B200SD_LAUNCH("softmax_rows_kernel", softmax_rows_kernel, dim3((unsigned)p->rows), dim3(256), smem, as_stream(stream), 
        p->x, p->in_dtype, p->ld_in, p->y, p->out_dtype == B200SD_F16, p->ld_out, p->cols, p->scale);
    count_launch();
    B200SD_CHECK_LAUNCH("softmax_rows");
    return B200SD_OK;
}

extern "C" int32_t b200sd_denorm_u8(const b200sd_denorm_u8_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->out, "denorm_u8: null pointer");
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype), "denorm_u8: bad in_dtype");
    B200SD_CHECK_ARG(p->n > 0 && p->n % 4 == 0 && ALIGNED16(p->x) && (reinterpret_cast<uintptr_t>(p->out) & 3) == 0,
                     "denorm_u8: n %% 4 == 0 and aligned pointers required");
    B200SD_LAUNCH("denorm_u8_kernel", denorm_u8_kernel, dim3(ew_grid(p->n / 4, 256)), dim3(256), 0, as_stream(stream), p->x, p->in_dtype, p->out, p->n);
    count_launch();
    B200SD_CHECK_LAUNCH("denorm_u8");
    return B200SD_OK;
}

extern "C" int32_t b200sd_cast(const b200sd_cast_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->x && p->y, "cast: null pointer");
    B200SD_CHECK_ARG(ANY_DTYPE(p->in_dtype) && ANY_DTYPE(p->out_dtype), "cast: bad dtype");
    B200SD_CHECK_ARG(p->n > 0 && p->n % 4 == 0 && ALIGNED16(p->x) && ALIGNED16(p->y), "cast: n %% 4 == 0, aligned");
    B200SD_LAUNCH("cast_kernel", cast_kernel, dim3(ew_grid(p->n / 4, 256)), dim3(256), 0, as_stream(stream), p->x, p->in_dtype, p->y, p->out_dtype, p->n);
    count_launch();
    B200SD_CHECK_LAUNCH("cast");
    return B200SD_OK;
}

// =============================================================================================== GroupNorm, no-allocation variant
// Some hosts that bind this library cannot hand it scratch memory.  A cgo / JNI / ctypes front end that only
// receives device pointers for its tensors from another component has no CUDA runtime of its own to allocate
// with, and a serving process that gave its whole device budget to a tensor arena at start-up has nothing left to
// carve a workspace from.  For them the split statistics of GroupNorm live in this array, which is part of the
// library image: every GPU that loads libb200sd.so holds these 256 KiB until the process exits.
#include <mutex>

namespace b200sd {

constexpr long long GN_STATIC_SCRATCH_FLOATS = 65536;  // 256 KiB; n_img * splits * groups * 2 must fit
static __device__ __align__(256) float g_gn_static_scratch[GN_STATIC_SCRATCH_FLOATS];

// In host code `g_gn_static_scratch` is only a handle for the symbol, not an address a kernel may dereference.
// The real device address is looked up once per device (with that device current) and cached.
static int32_t gn_static_scratch_address(float** out) {
    static std::mutex mu;
    static float* cached[64] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "groupnorm_static_scratch: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return set_error(B200SD_ERR_DEVICE, "groupnorm_static_scratch: device index %d not in [0, 64)", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (cached[dev] == nullptr) {
        float* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_static_scratch);
        if (e != cudaSuccess || p == nullptr)
            return set_error(B200SD_ERR_CUDA, "groupnorm_static_scratch: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        cached[dev] = p;
    }
    *out = cached[dev];
    return B200SD_OK;
}

}  // namespace b200sd

// Same computation and same two kernels as b200sd_groupnorm; `workspace` / `workspace_bytes` of the params are ignored.
//  * Allocates nothing: no cudaMalloc / cudaFree, no allocating library call.
//  * Limit: the split statistics, n_img * splits * groups * 2 floats, must fit the 65536 floats (256 KiB) of the array.
//    This entry point lowers the split count (down to 1) until they do, so it only refuses a problem when even one
//    split per image is too much: n_img * groups > 32768 (more than 1024 images at 32 groups).  Fewer splits mean fewer,
//    larger CTAs in the statistics pass: a large batch (n_img = 32 at 96x96) runs somewhat slower here than through
//    b200sd_groupnorm, with the same result up to the summation order of the split statistics.
//  * The first call on each device resolves the array's device address (cudaGetSymbolAddress); make that first call
//    before starting any stream capture on that device.
//  * All calls on a device share the one array and it keeps whatever the previous call left there.  Nothing has to be
//    cleared: the statistics kernel overwrites every slot the normalising kernel later reads.  Two calls whose
//    kernels overlap on the GPU (different streams, no dependency between them) race on it -- serialise them.
extern "C" int32_t b200sd_groupnorm_static_scratch(const b200sd_groupnorm_params* p, b200sd_stream_t stream) {
    int32_t rc = gn_check(p, "groupnorm_static_scratch");
    if (rc != B200SD_OK) return rc;
    B200SD_CHECK_ARG(!p->sums && !p->sums_out && p->y, "groupnorm_static_scratch: the split-phase fields (sums, sums_out) "
                     "belong to b200sd_groupnorm; this entry point always runs the whole GroupNorm");
    int splits = gn_splits(p->n_img, p->hw);
    const long long per_split = (long long)p->n_img * p->groups * 2;
    if (per_split > GN_STATIC_SCRATCH_FLOATS)
        return set_error(B200SD_ERR_WORKSPACE,
                         "groupnorm_static_scratch: needs %lld floats of scratch (n_img=%d x groups=%d x 2, one split) but the "
                         "built-in array holds %lld; use b200sd_groupnorm with a caller-owned workspace",
                         per_split, p->n_img, p->groups, GN_STATIC_SCRATCH_FLOATS);
    if ((long long)splits * per_split > GN_STATIC_SCRATCH_FLOATS) splits = (int)(GN_STATIC_SCRATCH_FLOATS / per_split);
    float* scratch = nullptr;
    rc = gn_static_scratch_address(&scratch);
    if (rc != B200SD_OK) return rc;
    return gn_launch(p, splits, scratch, as_stream(stream), "groupnorm_static_scratch");
}
