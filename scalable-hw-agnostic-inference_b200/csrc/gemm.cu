// K1/K2: persistent, warp-specialised implicit-GEMM for conv3x3 / conv1x1 / linear on sm_100a.
//
//   warp 0      TMA producer   : A = 4-D NHWC box (64 ch x bw x bh x bn pixels, shifted by the filter tap,
//                                zero-filled outside the image = the conv padding), B = 64 x BLOCK_N weight box
//   warp 1      UMMA issuer    : tcgen05.mma 128 x BLOCK_N x 16, fp32 accumulators in TMEM, double-buffered
//   warps 2..9  epilogue       : tcgen05.ld -> bias / timestep-row bias / residual / GEGLU -> 16-bit | fp32.  Two warps
//                                per TMEM lane quadrant take alternate 32-column chunks.  Linear layers without a
//                                residual leave by TMA store from a per-warp staging tile; everything else that has a
//                                plain column bias goes through the same tile transposed, so that global loads and
//                                stores are whole row segments; the rest (row bias, bias along M, ragged last
//                                chunk) stores straight from the accumulator row each thread holds.
//
// One CTA per SM; work items are walked n-fastest so an A tile is fetched from HBM once and the weight matrix
// stays L2-resident.  When a problem has fewer output tiles than SMs (the 24x24 and 12x12 UNet levels at small
// batch) the K loop is split across CTAs: every split writes its fp32 partial tile to a caller-owned workspace,
// and the CTA that arrives last at the tile's counter adds the partials in split order (deterministic) and runs
// the epilogue.  Algorithmic work per launch: 2*M*n_out*taps*cin FLOP.
#include "common.h"
#include "ptx.cuh"

namespace b200sd {
using namespace ptx;

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;
constexpr int A_TILE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KiB
constexpr int ACC_STRIDE = 256;                      // TMEM columns between the two accumulator buffers
constexpr int TMEM_COLS = 512;
constexpr int EPI_WARPS = 8;  // two warps per TMEM lane quadrant, interleaved over the 32-column chunks
constexpr int GEMM_THREADS = 64 + 32 * EPI_WARPS;
constexpr int MAX_SPLITS = 8;
constexpr int AUTO_SPLITS = 3;          // automatic K split factor (see plan_tiles)
constexpr int AUTO_SPLIT_MIN_KB = 128;  // ... only for K >= 8192
constexpr int AUTO_SPLIT_MAX_M_TILES = 4;  // ... and at most 512 output rows
constexpr int AUTO_FILL_MIN_KB = 64;       // second automatic rule: fill the machine when tiles <= SMs / 2 and K >= 4096
constexpr int WS_COUNTERS = 1024;  // int32 tile counters at the start of the split-K workspace

struct GemmKernelParams {
    int n_img, img_h, img_w, cin, taps;  // img_h / img_w: the OUTPUT feature map (= the input's for stride 1)
    int conv_stride, pad_lo;             // input pixel of output pixel x and tap s: x * conv_stride + s - pad_lo
    int a_bf16, b_bf16;  // operand formats for the instruction descriptor (0 = fp16, 1 = bf16)
    int bw, bh, bn;
    int tiles_x, tiles_y, m_tiles, n_tiles, kb_per_tap;
    int n_out, n_cols;  // n_cols = columns actually written (n_out, or n_out/2 for GEGLU)
    uint32_t a_box_bytes;
    int splits;          // >= 1
    float* partials;     // [tiles * splits][128][BLOCK_N] fp32 when splits > 1
    int* counters;       // [tiles] int32, zero between launches
    const float* bias;
    int bias_along_m;
    const float* row_bias;
    long long row_bias_stride;
    const void* residual;
    int res_dtype;
    long long ld_res;
    void* out;
    int out_dtype;
    long long ld_out;
    int geglu;
    int act;       // 1: erf-GELU on (acc + bias), before the residual
    int fast_epi;  // 1: outputs leave through the per-warp staging tile as full 128-byte row segments
    int tma_reduce;  // 1: (with tma_epi) out IS the residual (same dtype): the chunk is added to it by a TMA reduce, no residual read
    int tma_epi;   // 1: (linear layout, no residual) the staging tile is written in the output dtype, row per thread, and
                   //    leaves with one TMA store per 32x32 chunk: no shared-memory read-back on the epilogue's chain
    int res_prefetch;   // 1: the producer warp asks the tile's residual rows into L2 ahead of the epilogue (one bulk prefetch
                        //    per row, issued lane by lane)
    float2* tile_sums;  // (staged epilogue only) [m_tile * 4 + lane quadrant][n_out] (sum, sum of squares) of the stored
                        // values over the quadrant's valid rows: the next GroupNorm's statistics, or nullptr
};

// PAIR: two CTAs (one cluster) share a 256-row tile through tcgen05.mma.cta_group::2; each CTA stages its own 128 rows
// of A and only HALF of the weight tile, so the per-SM L2 traffic per k-block drops from (128 + N) to (128 + N/2) rows
// of 128 bytes.  Measured before: every tile width moved 55-70 B/clk/SM from L2 whatever the shape -- the operand
// fetch, not the tensor pipe, set the speed of the big convolutions.
template <int BLOCK_N, bool PAIR = false>
struct GemmCfg {
    static constexpr int B_TILE_BYTES = (PAIR ? BLOCK_N / 2 : BLOCK_N) * 128;
    static constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;
    static constexpr int STAGES = (200 * 1024) / STAGE_BYTES > 8 ? 8 : (200 * 1024) / STAGE_BYTES;
    // operand ring | barriers | per-tile column bias (2 x 256 fp32) | one 32x32 fp32 staging tile per epilogue warp.
    // No alignment slack: the kernel traps if the dynamic segment is not 1024-byte aligned (it always is in practice:
    // there is no static shared memory in front of it).
    static constexpr int STG_OFF = STAGES * STAGE_BYTES;  // 1024-byte aligned: the staging tiles are TMA-store sources
    static constexpr int BAR_OFF = STG_OFF + EPI_WARPS * 4096;
    static constexpr int BIAS_OFF = BAR_OFF + 256;
    static constexpr int SMEM_BYTES = BIAS_OFF + 2 * 256 * 4;
    static_assert(STG_OFF % 1024 == 0, "staging tiles must keep the swizzle-atom alignment");
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

// x * Phi(x), the erf-form GELU.  Phi(-u) = 2^Q(u) with Q a degree-7 fit of log2 Phi(-u) on [0, 6] (|dQ| < 6e-6, so
// Phi keeps ~4e-6 RELATIVE accuracy in the negative tail, where 0.5 * (1 + erff) cancels); Phi(u) = 1 - Phi(-u).
// 14 instructions per value instead of erff's 34: the GEGLU epilogue is instruction-bound.
__device__ __forceinline__ float gelu_erf(float x) {
    const float u = fminf(fabsf(x), 6.0f);
    float q = -1.889626233e-06f;
    q = fmaf(q, u, 6.268140101e-05f);
    q = fmaf(q, u, -9.388679333e-04f);
    q = fmaf(q, u, 8.539461332e-03f);
    q = fmaf(q, u, -5.402068712e-02f);
    q = fmaf(q, u, -4.584097768e-01f);
    q = fmaf(q, u, -1.151269147e+00f);
    q = fmaf(q, u, -9.999943403e-01f);
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(q));
    return x * (x < 0.f ? e : 1.0f - e);
}

// GEGLU on two (value, gate) pairs at once, polynomial and products on packed fp32 (FFMA2 / FMUL2 / FADD2).
// x * Phi(x) is formed as relu(x) - u * Phi(-u), u = min(|x|, 6): the same value as x * (x < 0 ? e : 1 - e) up to
// |x| * Phi(-6) ~ 1e-9 |x|, without the compare + predicated subtract per element.  Same polynomial as gelu_erf.
__device__ __forceinline__ void geglu_pair(uint64_t v2, uint64_t g2, float& o0, float& o1) {
    float g0, g1;
    unpack2(g2, g0, g1);
    const uint64_t u2 = pack2(fminf(fabsf(g0), 6.0f), fminf(fabsf(g1), 6.0f));
    uint64_t q2 = pack2(-1.889626233e-06f, -1.889626233e-06f);
    q2 = fma2(q2, u2, pack2(6.268140101e-05f, 6.268140101e-05f));
    q2 = fma2(q2, u2, pack2(-9.388679333e-04f, -9.388679333e-04f));
    q2 = fma2(q2, u2, pack2(8.539461332e-03f, 8.539461332e-03f));
    q2 = fma2(q2, u2, pack2(-5.402068712e-02f, -5.402068712e-02f));
    q2 = fma2(q2, u2, pack2(-4.584097768e-01f, -4.584097768e-01f));
    q2 = fma2(q2, u2, pack2(-1.151269147e+00f, -1.151269147e+00f));
    q2 = fma2(q2, u2, pack2(-9.999943403e-01f, -9.999943403e-01f));
    float q0, q1, e0, e1;
    unpack2(q2, q0, q1);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(q0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(q1));
    // relu(g) - u * e, then times the value
    const uint64_t gphi2 = fma2(u2, pack2(-e0, -e1), pack2(fmaxf(g0, 0.f), fmaxf(g1, 0.f)));
    unpack2(mul2(gphi2, v2), o0, o1);
}

// Applies the epilogue to 32 consecutive output columns [col0, col0+32) of one row and stores them.
__device__ __forceinline__ void epilogue_store32(const GemmKernelParams& p, float (&v)[32], long long m, int img,
                                                 int col0) {
    const bool full = (col0 + 32 <= p.n_cols);
    if (full) {
        if (p.residual) {
            if (p.res_dtype == B200SD_F32) {
                const float4* r = reinterpret_cast<const float4*>(static_cast<const float*>(p.residual) +
                                                                  m * p.ld_res + col0);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float4 t = r[j];
                    v[4 * j + 0] += t.x, v[4 * j + 1] += t.y, v[4 * j + 2] += t.z, v[4 * j + 3] += t.w;
                }
            } else {
                const bool f16 = p.res_dtype == B200SD_F16;
                const uint4* r = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.residual) +
                                                                m * p.ld_res + col0);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint4 t = r[j];
                    v[8 * j + 0] += h_lo(t.x, f16), v[8 * j + 1] += h_hi(t.x, f16);
                    v[8 * j + 2] += h_lo(t.y, f16), v[8 * j + 3] += h_hi(t.y, f16);
                    v[8 * j + 4] += h_lo(t.z, f16), v[8 * j + 5] += h_hi(t.z, f16);
                    v[8 * j + 6] += h_lo(t.w, f16), v[8 * j + 7] += h_hi(t.w, f16);
                }
            }
        }
        if (p.out_dtype == B200SD_F32) {
            float4* o = reinterpret_cast<float4*>(static_cast<float*>(p.out) + m * p.ld_out + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        } else {
            const bool f16 = p.out_dtype == B200SD_F16;
            uint4* o = reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out) + m * p.ld_out + col0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                o[j] = make_uint4(pack_h2(v[8 * j], v[8 * j + 1], f16), pack_h2(v[8 * j + 2], v[8 * j + 3], f16),
                                  pack_h2(v[8 * j + 4], v[8 * j + 5], f16), pack_h2(v[8 * j + 6], v[8 * j + 7], f16));
        }
    } else {
        for (int j = 0; j < 32; ++j) {
            int col = col0 + j;
            if (col >= p.n_cols) break;
            float x = v[j];
            if (p.residual) {
                if (p.res_dtype == B200SD_F32)
                    x += static_cast<const float*>(p.residual)[m * p.ld_res + col];
                else if (p.res_dtype == B200SD_F16)
                    x += __half2float(static_cast<const __half*>(p.residual)[m * p.ld_res + col]);
                else
                    x += __bfloat162float(static_cast<const __nv_bfloat16*>(p.residual)[m * p.ld_res + col]);
            }
            if (p.out_dtype == B200SD_F32)
                static_cast<float*>(p.out)[m * p.ld_out + col] = x;
            else if (p.out_dtype == B200SD_F16)
                static_cast<__half*>(p.out)[m * p.ld_out + col] = __float2half_rn(x);
            else
                static_cast<__nv_bfloat16*>(p.out)[m * p.ld_out + col] = __float2bfloat16(x);
        }
    }
}

// ---- staged epilogue -------------------------------------------------------------------------------------------
// tcgen05.ld hands every thread one accumulator ROW, so direct stores touch 32 different 128-byte lines per
// instruction (32 L1 wavefronts; measured: the fp32-residual linears ran at 65% LSU wavefront utilisation and 3 TB/s).
// Each epilogue warp therefore owns a 32x32 fp32 staging tile: rows are written as they come from TMEM (16-byte
// chunks XOR-swizzled by row, conflict-free), and read back transposed so that 8 (fp32) or 4 (16-bit) neighbouring
// lanes cover one row's 128 / 64 contiguous bytes.  Residual loads use the same mapping.
__device__ __forceinline__ void stage_rows(float* stg, int lane, const float (&v)[32]) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
__device__ __forceinline__ float4 stage_read(const float* stg, int row, int chunk) {
    return *reinterpret_cast<const float4*>(stg + row * 32 + ((chunk ^ (row & 7)) << 2));
}
// TMA-store variant: the thread writes its own row in the OUTPUT dtype, in the layout the store's tensor map expects
// (32 fp32 = 128-byte rows, SWIZZLE_128B: chunk ^= row & 7; 32 halves = 64-byte rows, SWIZZLE_64B: chunk ^= (row >> 1) & 3).
__device__ __forceinline__ void stage_rows_out(const GemmKernelParams& p, float* stg, int lane, const float (&v)[32]) {
    if (p.out_dtype == B200SD_F32) {
        stage_rows(stg, lane, v);
    } else {
        uint8_t* row = reinterpret_cast<uint8_t*>(stg) + lane * 64;
        const int sw = (lane >> 1) & 3;
        // one uniform branch around the sixteen conversions: chosen per conversion, the compiler issues BOTH encodings
        // under opposite predicates (32 issue slots instead of 16 per chunk)
        if (p.out_dtype == B200SD_F16) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4*>(row + ((j ^ sw) << 4)) =
                    make_uint4(pack_f16x2(v[8 * j], v[8 * j + 1]), pack_f16x2(v[8 * j + 2], v[8 * j + 3]),
                               pack_f16x2(v[8 * j + 4], v[8 * j + 5]), pack_f16x2(v[8 * j + 6], v[8 * j + 7]));
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4*>(row + ((j ^ sw) << 4)) =
                    make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                               pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        }
    }
}
// Transposed lane mapping.  fp32 outputs: 8 passes, row = 4*i + lane/8, columns 4*(lane%8)..+3.
//                            16-bit outputs: 4 passes, row = 8*i + lane/4, columns 8*(lane%4)..+7.
__device__ __forceinline__ void res_request(const GemmKernelParams& p, const int (&mt)[8], uint32_t vmask, int lane,
                                            int col0, uint4 (&buf)[8]) {
    if (p.out_dtype == B200SD_F32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = 4 * i + (lane >> 3);
            buf[i] = make_uint4(0, 0, 0, 0);
            if ((vmask >> row) & 1)
                buf[i] = *reinterpret_cast<const uint4*>(static_cast<const float*>(p.residual) +
                                                         (long long)mt[i] * p.ld_res + col0 + 4 * (lane & 7));
        }
    } else if (p.res_dtype == B200SD_F32) {
        // 16-bit output, fp32 residual (the feed-forward's second linear): 8 fp32 per lane and pass
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = 8 * i + (lane >> 2);
            buf[2 * i] = buf[2 * i + 1] = make_uint4(0, 0, 0, 0);
            if ((vmask >> row) & 1) {
                const uint4* src = reinterpret_cast<const uint4*>(static_cast<const float*>(p.residual) +
                                                                  (long long)mt[i] * p.ld_res + col0 + 8 * (lane & 3));
                buf[2 * i] = src[0], buf[2 * i + 1] = src[1];
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = 8 * i + (lane >> 2);
            buf[i] = make_uint4(0, 0, 0, 0);
            if ((vmask >> row) & 1)
                buf[i] = *reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.residual) +
                                                         (long long)mt[i] * p.ld_res + col0 + 8 * (lane & 3));
        }
    }
}
// Reads the staged chunk back transposed, adds the column bias (`bs` = this chunk's 32 bias values in shared memory,
// or nullptr) and the requested residual (`use_res`), converts and stores full row segments.
// `sums` (or nullptr) = where this warp's column sums of the chunk go: after the stores, the per-lane partial sums over
// the rows a lane handled are combined across the lanes that share its columns (xor butterfly: every lane adds the
// same values in the same order, so the result does not depend on timing) and one lane per column group writes them.
__device__ __forceinline__ void staged_store(const GemmKernelParams& p, const float* stg, const int (&mt)[8],
                                             uint32_t vmask, int lane, int col0, const float* bs, bool use_res,
                                             const uint4 (&buf)[8], float2* sums) {
    if (p.out_dtype == B200SD_F32) {
        const int cg = lane & 7;
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bs) b = *reinterpret_cast<const float4*>(bs + 4 * cg);
        float4 cs = make_float4(0.f, 0.f, 0.f, 0.f), cq = cs;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = 4 * i + (lane >> 3);
            float4 a = stage_read(stg, row, cg);
            a.x += b.x, a.y += b.y, a.z += b.z, a.w += b.w;
            if (p.act) a.x = gelu_erf(a.x), a.y = gelu_erf(a.y), a.z = gelu_erf(a.z), a.w = gelu_erf(a.w);
            if (use_res) {
                a.x += __uint_as_float(buf[i].x), a.y += __uint_as_float(buf[i].y);
                a.z += __uint_as_float(buf[i].z), a.w += __uint_as_float(buf[i].w);
            }
            if ((vmask >> row) & 1) {
                *reinterpret_cast<float4*>(static_cast<float*>(p.out) + (long long)mt[i] * p.ld_out + col0 + 4 * cg) = a;
                if (sums) {
                    cs.x += a.x, cs.y += a.y, cs.z += a.z, cs.w += a.w;
                    cq.x = fmaf(a.x, a.x, cq.x), cq.y = fmaf(a.y, a.y, cq.y);
                    cq.z = fmaf(a.z, a.z, cq.z), cq.w = fmaf(a.w, a.w, cq.w);
                }
            }
        }
        if (sums) {
#pragma unroll
            for (int o = 8; o <= 16; o <<= 1) {
                cs.x += __shfl_xor_sync(0xffffffffu, cs.x, o), cs.y += __shfl_xor_sync(0xffffffffu, cs.y, o);
                cs.z += __shfl_xor_sync(0xffffffffu, cs.z, o), cs.w += __shfl_xor_sync(0xffffffffu, cs.w, o);
                cq.x += __shfl_xor_sync(0xffffffffu, cq.x, o), cq.y += __shfl_xor_sync(0xffffffffu, cq.y, o);
                cq.z += __shfl_xor_sync(0xffffffffu, cq.z, o), cq.w += __shfl_xor_sync(0xffffffffu, cq.w, o);
            }
            if (lane < 8) {
                float4* d = reinterpret_cast<float4*>(sums + col0 + 4 * cg);
                d[0] = make_float4(cs.x, cq.x, cs.y, cq.y);
                d[1] = make_float4(cs.z, cq.z, cs.w, cq.w);
            }
        }
    } else {
        const bool f16 = p.out_dtype == B200SD_F16;
        const int cg = lane & 3;
        float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
        if (bs) {
            b0 = *reinterpret_cast<const float4*>(bs + 8 * cg);
            b1 = *reinterpret_cast<const float4*>(bs + 8 * cg + 4);
        }
        float cs[8], cq[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) cs[j] = 0.f, cq[j] = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = 8 * i + (lane >> 2);
            float4 a0 = stage_read(stg, row, 2 * cg), a1 = stage_read(stg, row, 2 * cg + 1);
            a0.x += b0.x, a0.y += b0.y, a0.z += b0.z, a0.w += b0.w;
            a1.x += b1.x, a1.y += b1.y, a1.z += b1.z, a1.w += b1.w;
            if (p.act) {
                a0.x = gelu_erf(a0.x), a0.y = gelu_erf(a0.y), a0.z = gelu_erf(a0.z), a0.w = gelu_erf(a0.w);
                a1.x = gelu_erf(a1.x), a1.y = gelu_erf(a1.y), a1.z = gelu_erf(a1.z), a1.w = gelu_erf(a1.w);
            }
            if (use_res && p.res_dtype == B200SD_F32) {
                a0.x += __uint_as_float(buf[2 * i].x), a0.y += __uint_as_float(buf[2 * i].y);
                a0.z += __uint_as_float(buf[2 * i].z), a0.w += __uint_as_float(buf[2 * i].w);
                a1.x += __uint_as_float(buf[2 * i + 1].x), a1.y += __uint_as_float(buf[2 * i + 1].y);
                a1.z += __uint_as_float(buf[2 * i + 1].z), a1.w += __uint_as_float(buf[2 * i + 1].w);
            } else if (use_res) {
                a0.x += h_lo(buf[i].x, f16), a0.y += h_hi(buf[i].x, f16);
                a0.z += h_lo(buf[i].y, f16), a0.w += h_hi(buf[i].y, f16);
                a1.x += h_lo(buf[i].z, f16), a1.y += h_hi(buf[i].z, f16);
                a1.z += h_lo(buf[i].w, f16), a1.w += h_hi(buf[i].w, f16);
            }
            if ((vmask >> row) & 1) {
                *reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out) + (long long)mt[i] * p.ld_out + col0 + 8 * cg) =
                    make_uint4(pack_h2(a0.x, a0.y, f16), pack_h2(a0.z, a0.w, f16), pack_h2(a1.x, a1.y, f16),
                               pack_h2(a1.z, a1.w, f16));
                if (sums) {
                    const float v[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) cs[j] += v[j], cq[j] = fmaf(v[j], v[j], cq[j]);
                }
            }
        }
        if (sums) {
#pragma unroll
            for (int o = 4; o <= 16; o <<= 1)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], o);
                    cq[j] += __shfl_xor_sync(0xffffffffu, cq[j], o);
                }
            if (lane < 4) {
                float4* d = reinterpret_cast<float4*>(sums + col0 + 8 * cg);
#pragma unroll
                for (int j = 0; j < 4; ++j) d[j] = make_float4(cs[2 * j], cq[2 * j], cs[2 * j + 1], cq[2 * j + 1]);
            }
        }
    }
}

// adds bias[n0 + j] (n0 = index into the bias vector, guarded by n_out) and the per-image row bias
__device__ __forceinline__ void add_col_bias32(const GemmKernelParams& p, float (&v)[32], long long m, int img,
                                               int n0) {
    if (p.bias) {
        if (p.bias_along_m) {
            float b = p.bias[m];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += b;
        } else if (n0 + 32 <= p.n_out) {
            const float4* b = reinterpret_cast<const float4*>(p.bias + n0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4 t = __ldg(b + j);
                v[4 * j + 0] += t.x, v[4 * j + 1] += t.y, v[4 * j + 2] += t.z, v[4 * j + 3] += t.w;
            }
        } else {
            for (int j = 0; j < 32 && n0 + j < p.n_out; ++j) v[j] += p.bias[n0 + j];
        }
    }
    if (p.row_bias) {
        const float* rb = p.row_bias + (long long)img * p.row_bias_stride + n0;
        if (n0 + 32 <= p.n_out) {
            const float4* b = reinterpret_cast<const float4*>(rb);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4 t = __ldg(b + j);
                v[4 * j + 0] += t.x, v[4 * j + 1] += t.y, v[4 * j + 2] += t.z, v[4 * j + 3] += t.w;
            }
        } else {
            for (int j = 0; j < 32 && n0 + j < p.n_out; ++j) v[j] += rb[j];
        }
    }
}

template <int BLOCK_N, bool PAIR>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmOut, const GemmKernelParams p) {
    using Cfg = GemmCfg<BLOCK_N, PAIR>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // SWIZZLE_128B tiles need 1024-byte alignment in the shared address space
    uint8_t* smem = smem_raw;
    if (smem_u32(smem_raw) & 1023u) __trap();
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::BAR_OFF);
    float* bias_s = reinterpret_cast<float*>(smem + Cfg::BIAS_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    int* last_flag = reinterpret_cast<int*>(tmem_base_slot + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], PAIR ? 2 * EPI_WARPS : EPI_WARPS);  // pair: both CTAs' epilogues report to the leader
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        if (PAIR) tmem_alloc_pair(tmem_base_slot, TMEM_COLS);
        else tmem_alloc(tmem_base_slot, TMEM_COLS);
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all();  // the peer's barriers exist before anything is signalled across the pair
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    pdl_sync();  // everything above overlapped the previous kernel's tail; global memory is touched from here on

    const int splits = p.splits;
    const int num_kb = p.taps * p.kb_per_tap;
    // Work items.  Single CTA: w = tile * splits + split, tile = m_tile * n_tiles + n_tile.  Pair: w = pair tile, the
    // two CTAs take m_tile = 2 * (w / n_tiles) + rank of the same n_tile.
    const int rank = PAIR ? (int)cluster_ctarank() : 0;
    const int w_first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int w_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int total_work = PAIR ? ((p.m_tiles + 1) / 2) * p.n_tiles : p.m_tiles * p.n_tiles * splits;
    auto m_tile_of = [&](int t) { return PAIR ? 2 * (t / p.n_tiles) + rank : t / p.n_tiles; };
    // columns the UMMA of n_tile computes (pair mode narrows the last tile instead of multiplying zeros)
    auto n_this_of = [&](int n_tile) {
        const int left = p.n_out - n_tile * BLOCK_N;
        return left >= BLOCK_N ? BLOCK_N : ((left + 31) / 32) * 32;
    };

    // Warps 0 and 1 keep all 32 lanes in the (warp-uniform) control flow and predicate only the asynchronous
    // instructions on one elected lane.  Run from inside an `if (lane == 0)` region instead, every TMA / UMMA operand
    // would count as thread-varying and the compiler would wrap each instruction in an R2UR + ELECT + BRA.U.ANY
    // waterfall loop (measured in the attention kernel: ~5 cycles per scalar instruction on the single issuing
    // thread, which made instruction ISSUE, not the tensor pipe, the limit).
    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        int stage = 0;
        uint32_t phase = 0;
        const int res_esz = p.res_dtype == B200SD_F32 ? 4 : 2;
        const bool res_l2_prefetch =
            p.res_prefetch && p.residual != nullptr && splits == 1 && !p.geglu && !p.tma_reduce &&
            ((reinterpret_cast<uintptr_t>(p.residual) | (uintptr_t)(p.ld_res * res_esz)) & 15) == 0;
        for (int w = w_first; w < total_work; w += w_step) {
            const int t = w / splits, split = w - t * splits;
            const int n_tile = t % p.n_tiles;
            const int m_tile = m_tile_of(t);
            const int tx = m_tile % p.tiles_x;
            const int ty = (m_tile / p.tiles_x) % p.tiles_y;
            const int tn = m_tile / (p.tiles_x * p.tiles_y);
            const int x0 = tx * p.bw, y0 = ty * p.bh, n0 = tn * p.bn;
            const int kb0 = (int)((long long)split * num_kb / splits);
            const int kb1 = (int)((long long)(split + 1) * num_kb / splits);
            if (res_l2_prefetch) {
                // The residual rows of this tile are asked into L2 now, a pipeline depth ahead of the epilogue that
                // adds them: its row-per-thread loads then see L2 latency instead of HBM latency.
                const int colb = n_tile * BLOCK_N;
                const int ncol = (p.n_cols - colb) < BLOCK_N ? (p.n_cols - colb) : BLOCK_N;
                const uint32_t bytes = (uint32_t)(ncol * res_esz);
                if ((bytes & 15u) == 0) {
                    const int box_pix = p.bw * p.bh;
                    for (int r = lane; r < BLOCK_M; r += 32) {
                        const int ln = r / box_pix;
                        const int rem = r - ln * box_pix;
                        const int ly = rem / p.bw;
                        const int img = n0 + ln, y = y0 + ly, x = x0 + (rem - ly * p.bw);
                        if (ln < p.bn && img < p.n_img && y < p.img_h && x < p.img_w) {
                            const long long m = ((long long)img * p.img_h + y) * p.img_w + x;
                            l2_prefetch_bulk(static_cast<const uint8_t*>(p.residual) +
                                                 (m * p.ld_res + colb) * (long long)res_esz, bytes);
                        }
                    }
                }
                __syncwarp();
            }
            for (int kb = kb0; kb < kb1; ++kb) {
                const int tap = kb / p.kb_per_tap;
                const int c0 = (kb - tap * p.kb_per_tap) * BLOCK_K;
                // first input pixel of the box for this tap: x0 * stride + s - pad_lo (stride 2: the tensor map's element
                // strides make the box take every second pixel from there)
                int dx = x0, dy = y0;
                if (p.taps == 9) {
                    dy = y0 * p.conv_stride + tap / 3 - p.pad_lo;
                    dx = x0 * p.conv_stride + tap - (tap / 3) * 3 - p.pad_lo;
                }
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
                uint8_t* sb = sa + A_TILE_BYTES;
                if (elect_one()) {
                    if (PAIR) {
                        // both CTAs' bytes are counted on the leader's barrier; each CTA brings its half of the weight rows
                        if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * (p.a_box_bytes + Cfg::B_TILE_BYTES));
                        tma_load_4d_pair(sa, &tmA, &full_bar[stage], c0, dx, dy, n0);
                        tma_load_2d_pair(sb, &tmB, &full_bar[stage], tap * p.cin + c0,
                                         n_tile * BLOCK_N + rank * (n_this_of(n_tile) >> 1));
                    } else {
                        mbar_expect_tx(&full_bar[stage], p.a_box_bytes + Cfg::B_TILE_BYTES);
                        tma_load_4d(sa, &tmA, &full_bar[stage], c0, dx, dy, n0);
                        tma_load_2d(sb, &tmB, &full_bar[stage], tap * p.cin + c0, n_tile * BLOCK_N);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ------------------------------------------------------------------ UMMA issuer (pair: the leader CTA only)
        uint32_t idesc = umma_idesc(BLOCK_M, BLOCK_N, (uint32_t)p.a_bf16, (uint32_t)p.b_bf16);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t smem_u = __shfl_sync(0xffffffffu, smem_u32(smem), 0);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int w = w_first; w < total_work; w += w_step) {
            const int split = w % splits;
            const int nkb = (int)((long long)(split + 1) * num_kb / splits) - (int)((long long)split * num_kb / splits);
            if (PAIR) idesc = umma_idesc(2 * BLOCK_M, n_this_of(w % p.n_tiles), (uint32_t)p.a_bf16, (uint32_t)p.b_bf16);
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_u + acc * ACC_STRIDE;
            for (int kb = 0; kb < nkb; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint32_t a_addr = smem_u + stage * Cfg::STAGE_BYTES;
                const uint64_t adesc = umma_desc_sw128(a_addr);
                const uint64_t bdesc = umma_desc_sw128(a_addr + A_TILE_BYTES);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < BLOCK_K / 16; ++k) {
                        // +32 bytes along K inside the 128-byte swizzle row = +2 in the (addr >> 4) field
                        if (PAIR) umma_bf16_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                        else umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    if (PAIR) {
                        // the stage is free, and the accumulator ready, in BOTH CTAs
                        umma_commit_pair(&empty_bar[stage], 0b11);
                        if (kb == nkb - 1) umma_commit_pair(&tmem_full[acc], 0b11);
                    } else {
                        umma_commit(&empty_bar[stage]);
                        if (kb == nkb - 1) umma_commit(&tmem_full[acc]);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    } else if (warp >= 2) {
        // ------------------------------------------------------------------ epilogue (warps 2..9)
        // The epilogue is bound by bytes in flight (residual reads) or by ALU work (GEGLU), not by TMEM bandwidth, so
        // two warps share each lane quadrant: warp group `eg` takes the 32-column chunks c = eg, eg + 2, ...
        const int quad = warp & 3;  // TMEM lane quadrant this warp may read
        const int eg = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;  // 0..255 over the epilogue warps
        float* stg = reinterpret_cast<float*>(smem + Cfg::STG_OFF) + (warp - 2) * 1024;
        // TMA-store epilogue with 16-bit outputs: a 32x32 chunk is 2 KiB, so the warp's 4 KiB staging region holds TWO
        // tiles and the warp stages chunk c + 1 while the store of chunk c still reads the other one (with one tile the
        // warp idled in bulk_wait_read for the whole store latency: 14 % of the samples of the level-0 QKV projection)
        const bool stg2 = p.tma_epi != 0 && p.out_dtype != B200SD_F32;
        int stg_sel = 0;
        const bool out32 = p.out_dtype == B200SD_F32;
        const bool fast = p.fast_epi != 0;
        const bool tma = p.tma_epi != 0;
        const int row_in_tile = quad * 32 + lane;
        const int box_pix = p.bw * p.bh;
        // position of this thread's accumulator row inside the pixel box (the same for every tile)
        const int ln = row_in_tile / box_pix;
        const int ly = (row_in_tile - ln * box_pix) / p.bw;
        const int lx = row_in_tile - ln * box_pix - ly * p.bw;
        // the column bias of a tile is fetched one tile ahead (one value per epilogue thread)
        const bool col_bias = p.bias != nullptr && !p.bias_along_m;
        // (a per-image row bias joins it when every tile lies inside one image -- the staged epilogue then needs no
        // per-row lookup; fast_epi is only set with a row bias if bn == 1)
        const bool row_bias_in_tile = fast && p.row_bias != nullptr;
        auto bias_fetch = [&](int w) -> float {
            const int t2 = w / splits;
            const int col = (t2 % p.n_tiles) * BLOCK_N + et;
            if (!(et < BLOCK_N && w < total_work && col < p.n_out)) return 0.f;
            float b = col_bias ? __ldg(p.bias + col) : 0.f;
            if (row_bias_in_tile) {
                const int img = m_tile_of(t2) / (p.tiles_x * p.tiles_y);
                if (img < p.n_img) b += __ldg(p.row_bias + (long long)img * p.row_bias_stride + col);
            }
            return b;
        };
        float bias_cur = bias_fetch(w_first);
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int w = w_first; w < total_work; w += w_step) {
            const int t = w / splits;
            const int n_tile = t % p.n_tiles;
            const int m_tile = m_tile_of(t);
            const int tx = m_tile % p.tiles_x;
            const int ty = (m_tile / p.tiles_x) % p.tiles_y;
            const int tn = m_tile / (p.tiles_x * p.tiles_y);
            const int img = tn * p.bn + ln;
            const int y = ty * p.bh + ly;
            const int x = tx * p.bw + lx;
            const bool valid = (ln < p.bn) && (img < p.n_img) && (y < p.img_h) && (x < p.img_w);
            const long long m = ((long long)img * p.img_h + y) * p.img_w + x;

            // column bias of this tile -> shared memory (read back as broadcasts); double-buffered with the accumulator
            float* bs = bias_s + acc * 256;
            if (et < BLOCK_N) bs[et] = bias_cur;
            bias_cur = bias_fetch(w + w_step);
            // rows this lane handles in the transposed (staged) mapping
            const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
            int mt[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                mt[i] = __shfl_sync(0xffffffffu, (int)m, out32 ? 4 * i + (lane >> 3) : 8 * (i & 3) + (lane >> 2));
            const int colbase = n_tile * BLOCK_N;
            const int row0 = m_tile * BLOCK_M + quad * 32;  // first output row of this warp (linear layout only)
            const bool use_res = fast && !tma && p.residual != nullptr;  // (a TMA reduce needs no residual read)
            // this warp's row of the column-sum table: slice (m_tile, lane quadrant); a pair's odd CTA may hold a tile
            // past the end of the problem
            float2* sums_row = (p.tile_sums != nullptr && m_tile < p.m_tiles)
                                   ? p.tile_sums + (size_t)(m_tile * 4 + quad) * p.n_out : nullptr;
            uint4 rbuf[8];
            if (use_res && splits == 1 && colbase + eg * 32 + 32 <= p.n_cols)
                res_request(p, mt, vmask, lane, colbase + eg * 32, rbuf);
            named_bar_sync(2, 32 * EPI_WARPS);  // bias tile complete

            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_addr = tmem_base + acc * ACC_STRIDE + ((uint32_t)(quad * 32) << 16);
            if (splits > 1) {
                // ---- split-K: park the partial tile, release TMEM, elect the last-arriving split to finish the tile
                float* mine = p.partials + ((size_t)w * BLOCK_M + row_in_tile) * BLOCK_N;
#pragma unroll 1
                for (int c = eg; c < BLOCK_N / 32; c += 2) {
                    uint32_t r[32];
                    tmem_ld_x32(t_addr + c * 32, r);
                    tmem_ld_wait();
                    float4* o = reinterpret_cast<float4*>(mine + c * 32);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        o[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                           __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { if (PAIR) mbar_arrive_leader(&tmem_empty[acc]); else mbar_arrive(&tmem_empty[acc]); }
                named_bar_sync(1, 32 * EPI_WARPS);  // all 128 rows of the partial tile are written ...
                if (row_in_tile == 0 && eg == 0) {
                    __threadfence();     // ... and (fence cumulativity) visible device-wide before the count goes up
                    const int old = atomicAdd(p.counters + t, 1);
                    const int last = (old == splits - 1);
                    if (last) {
                        p.counters[t] = 0;  // leave the workspace clean for the next launch
                        __threadfence();
                    }
                    *last_flag = last;
                }
                named_bar_sync(1, 32 * EPI_WARPS);
                const bool last = *last_flag != 0;
                named_bar_sync(1, 32 * EPI_WARPS);  // everyone has read the flag before the next work item rewrites it
                if (last) {
                    const float* base = p.partials + ((size_t)t * splits * BLOCK_M + row_in_tile) * BLOCK_N;
#pragma unroll 1
                    for (int c = eg; c < BLOCK_N / 32; c += 2) {
                        const int col0 = n_tile * BLOCK_N + c * 32;
                        if (valid && col0 < p.n_cols) {
                            float v[32];
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = 0.f;
                            for (int s = 0; s < splits; ++s) {
                                const float4* src =
                                    reinterpret_cast<const float4*>(base + (size_t)s * BLOCK_M * BLOCK_N + c * 32);
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const float4 q = __ldcg(src + j);
                                    v[4 * j] += q.x, v[4 * j + 1] += q.y, v[4 * j + 2] += q.z, v[4 * j + 3] += q.w;
                                }
                            }
                            add_col_bias32(p, v, m, img, col0);
                            if (p.act) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                            }
                            epilogue_store32(p, v, m, img, col0);
                        }
                    }
                }
            } else {
                // Chunk loops are NOT unrolled: the kernel's code size (instruction-cache misses were 11% of the
                // epilogue's stall samples) matters more than the loop overhead.  The TMEM load of a warp's next chunk
                // is issued as soon as the registers of the current one are staged, and the accumulator buffer is
                // handed back to the UMMA warp right after the last load, before the last stores.
                if (!p.geglu) {
                    constexpr int NC = BLOCK_N / 32;
                    uint32_t r[32];
                    if (eg < NC) tmem_ld_x32(t_addr + eg * 32, r);
#pragma unroll 1
                    for (int c = eg; c < NC; c += 2) {
                        const int col0 = colbase + c * 32;
                        const bool more = c + 2 < NC;
                        if (tma) {
                            // bias reads are issued before the TMEM wait so the two latencies overlap
                            float4 bq[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j) bq[j] = *reinterpret_cast<const float4*>(bs + c * 32 + 4 * j);
                            const bool live = col0 < p.n_cols && row0 < p.img_w;
                            tmem_ld_wait();
                            float v[32];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {  // bias adds on packed pairs (FADD2)
                                unpack2(add2(pack2(__uint_as_float(r[4 * j + 0]), __uint_as_float(r[4 * j + 1])), pack2(bq[j].x, bq[j].y)),
                                        v[4 * j + 0], v[4 * j + 1]);
                                unpack2(add2(pack2(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])), pack2(bq[j].z, bq[j].w)),
                                        v[4 * j + 2], v[4 * j + 3]);
                            }
                            if (p.act) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                            }
                            // (the tiles alternate per ISSUED store -- a chunk past the edge commits no bulk group, and
                            // wait_group.read 1 only vouches for the store before the most recent one)
                            float* stg_c = stg + (stg2 ? stg_sel * 512 : 0);
                            if (live) stg_sel ^= 1;
                            if (lane == 0) {  // the store that last read this staging tile let go of it
                                if (stg2) bulk_wait_read1();
                                else bulk_wait_read0();
                            }
                            __syncwarp();
                            stage_rows_out(p, stg_c, lane, v);
                            if (more) {
                                tmem_ld_x32(t_addr + (c + 2) * 32, r);
                            } else {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) { if (PAIR) mbar_arrive_leader(&tmem_empty[acc]); else mbar_arrive(&tmem_empty[acc]); }
                            }
                            fence_proxy_async_smem();
                            __syncwarp();
                            if (lane == 0 && live) {
                                // rows / columns past the edge are clipped; with tma_reduce the tile is added to the
                                // residual that already sits in `out`
                                if (p.tma_reduce) tma_reduce_add_2d(&tmOut, stg_c, col0, row0);
                                else tma_store_2d(&tmOut, stg_c, col0, row0);
                                bulk_commit();
                            }
                            continue;
                        }
                        tmem_ld_wait();
                        const bool full = fast && col0 + 32 <= p.n_cols;
                        if (full) {
                            float v[32];
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                            stage_rows(stg, lane, v);
                        } else if (valid && col0 < p.n_cols) {  // ragged last chunk, row bias, bias along M, mixed dtypes
                            float v[32];
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                            add_col_bias32(p, v, m, img, col0);
                            if (p.act) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                            }
                            epilogue_store32(p, v, m, img, col0);
                        }
                        if (more) {
                            tmem_ld_x32(t_addr + (c + 2) * 32, r);
                        } else {
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) { if (PAIR) mbar_arrive_leader(&tmem_empty[acc]); else mbar_arrive(&tmem_empty[acc]); }
                        }
                        __syncwarp();
                        if (full) {
                            staged_store(p, stg, mt, vmask, lane, col0, bs + c * 32, use_res, rbuf, sums_row);
                            if (more && use_res && col0 + 96 <= p.n_cols)
                                res_request(p, mt, vmask, lane, col0 + 64, rbuf);
                            __syncwarp();
                        }
                    }
                    if (eg >= NC) {  // BLOCK_N = 32 * odd: the second warp group may have no chunk at all
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) { if (PAIR) mbar_arrive_leader(&tmem_empty[acc]); else mbar_arrive(&tmem_empty[acc]); }
                    }
                } else {
                    constexpr int HALF = BLOCK_N / 2;
                    uint32_t rv[32], rg[32];
                    tmem_ld_x32(t_addr + eg * 32, rv);
                    tmem_ld_x32(t_addr + HALF + eg * 32, rg);
#pragma unroll 1
                    for (int c = eg; c < HALF / 32; c += 2) {
                        const bool more = c + 2 < HALF / 32;
                        tmem_ld_wait();
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 bv = *reinterpret_cast<const float4*>(bs + c * 32 + 4 * j);
                            const float4 bg = *reinterpret_cast<const float4*>(bs + HALF + c * 32 + 4 * j);
                            // bias adds on packed pairs as well (FADD2)
                            geglu_pair(add2(pack2(__uint_as_float(rv[4 * j + 0]), __uint_as_float(rv[4 * j + 1])), pack2(bv.x, bv.y)),
                                       add2(pack2(__uint_as_float(rg[4 * j + 0]), __uint_as_float(rg[4 * j + 1])), pack2(bg.x, bg.y)),
                                       v[4 * j + 0], v[4 * j + 1]);
                            geglu_pair(add2(pack2(__uint_as_float(rv[4 * j + 2]), __uint_as_float(rv[4 * j + 3])), pack2(bv.z, bv.w)),
                                       add2(pack2(__uint_as_float(rg[4 * j + 2]), __uint_as_float(rg[4 * j + 3])), pack2(bg.z, bg.w)),
                                       v[4 * j + 2], v[4 * j + 3]);
                        }
                        if (more) {
                            tmem_ld_x32(t_addr + (c + 2) * 32, rv);
                            tmem_ld_x32(t_addr + HALF + (c + 2) * 32, rg);
                        } else {
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) { if (PAIR) mbar_arrive_leader(&tmem_empty[acc]); else mbar_arrive(&tmem_empty[acc]); }
                        }
                        const int col0 = n_tile * HALF + c * 32;
                        if (tma) {
                            float* stg_c = stg + (stg2 ? stg_sel * 512 : 0);
                            if (row0 < p.img_w) stg_sel ^= 1;
                            if (lane == 0) {
                                if (stg2) bulk_wait_read1();
                                else bulk_wait_read0();
                            }
                            __syncwarp();
                            stage_rows_out(p, stg_c, lane, v);
                            fence_proxy_async_smem();
                            __syncwarp();
                            if (lane == 0 && row0 < p.img_w) {
                                tma_store_2d(&tmOut, stg_c, col0, row0);
                                bulk_commit();
                            }
                        } else if (fast) {
                            stage_rows(stg, lane, v);
                            __syncwarp();
                            staged_store(p, stg, mt, vmask, lane, col0, nullptr, false, rbuf, nullptr);
                            __syncwarp();
                        } else if (valid) {
                            epilogue_store32(p, v, m, img, col0);
                        }
                    }
                }
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        if (tma && lane == 0) bulk_wait0();  // the staging tiles must outlive the stores that read them
    }

    tc_fence_before();
    if (PAIR) cluster_sync_all();  // neither CTA frees TMEM or exits while the other may still signal it
    else __syncthreads();
    if (warp == 2) {
        if (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS);
        else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------ host side
static void choose_box(int n_img, int h, int w, int* bw, int* bh, int* bn) {
    if (h == 1 && n_img == 1) {
        *bw = w < 128 ? w : 128, *bh = 1, *bn = 1;
        return;
    }
    double best = -1.0;
    int b_w = 1, b_h = 1, b_n = 1;
    const int maxw = w < 128 ? w : 128;
    for (int cw = maxw; cw >= 1; --cw) {
        const int maxh = (128 / cw) < h ? (128 / cw) : h;
        for (int ch = maxh; ch >= 1; --ch) {
            // full-width boxes may take the same rows of several images (the 4-D TMA box is n x h x w x c, and a row
            // outside the image is outside the tensor for every image of the box, so the zero padding still holds):
            // a 12x12 image is 2 tiles at 56 % as 12x10 + 12x2, but 2 rows x 5 images fill 120 of 128 rows
            int cn = 1;
            if (cw == w) {
                cn = 128 / (cw * ch);
                if (cn > n_img) cn = n_img;
                if (cn < 1) cn = 1;
            }
            const long long tiles = (long long)((w + cw - 1) / cw) * ((h + ch - 1) / ch) * ((n_img + cn - 1) / cn);
            // efficiency, with a mild preference for long contiguous rows and for boxes inside one image
            const double eff = (double)w * h * n_img / (double)(tiles * 128) + 1e-4 * cw - (cn > 1 ? 1e-3 : 0.0);
            if (eff > best) best = eff, b_w = cw, b_h = ch, b_n = cn;
        }
    }
    *bw = b_w, *bh = b_h, *bn = b_n;
}

static int pick_block_n_auto(int n_out, int geglu, long long m_tiles) {
    if (geglu) return (n_out % 256 == 0) ? 256 : 128;
    if (n_out <= 64) return 64;
    // The main loop is bound by L2->SM bandwidth per SM (measured, tools/gemm_sweep.py): a work item costs
    // (16 KiB of A + 128 B x block_n of B) per k-block, and items run in ceil(tiles / SMs) waves.  Pick the width
    // that minimises waves x bytes; ties go to the wider tile.
    const long long sms = num_sms();
    const int cands[4] = {256, 160, 128, 64};
    int best = 128;
    long long best_cost = -1;
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        const long long tiles = m_tiles * ((n_out + bn - 1) / bn);
        const long long cost = ((tiles + sms - 1) / sms) * (128 + bn);
        if (best_cost < 0 || cost < best_cost) best_cost = cost, best = bn;
    }
    return best;
}

// Tile width and K-split for a problem.  With a workspace, under-filled problems keep the widest tile (least L2->SM
// traffic per FLOP) and split K instead of shrinking the tile.
static void plan_tiles(int n_out, int geglu, long long m_tiles, int num_kb, int user_block_n, int user_split,
                       bool have_ws, int* block_n, int* splits) {
    *splits = 1;
    if (user_block_n) {
        *block_n = user_block_n;
    } else {
        *block_n = pick_block_n_auto(n_out, geglu, m_tiles);
    }
    if (geglu || !have_ws || user_split == 1) return;
    const int sms = num_sms();
    long long tiles = m_tiles * ((n_out + *block_n - 1) / *block_n);
    int s = 1;
    if (user_split > 1) {
        s = user_split;
    } else if (!user_block_n && m_tiles <= AUTO_SPLIT_MAX_M_TILES && num_kb >= AUTO_SPLIT_MIN_KB) {
        // Measured (tools/gemm_sweep.py): parking partials + the counter handshake cost several us per work item, so
        // splitting pays only for long K loops over very few rows -- the 12x12 UNet level at small batch
        // (M = 288, K = 11520 / 23040): 128-wide tiles x 3 splits run 1.4-1.8x faster than any unsplit tiling.
        const long long t128 = m_tiles * ((n_out + 127) / 128);
        if (t128 * AUTO_SPLITS <= sms) {
            *block_n = 128;
            tiles = t128;
            s = AUTO_SPLITS;
        }
    }
    if (s == 1 && user_split == 0 && !user_block_n && num_kb >= AUTO_FILL_MIN_KB && tiles * 2 <= sms) {
        // Under-filled problems with a long K loop (the 24x24 level at CFG batch 2: 72 tiles of 180 k-blocks on 148 SMs):
        // as many splits as fit the machine.  Shorter loops stay whole -- parking and re-reading the partial tiles costs
        // more than the idle SMs.
        s = (int)(sms / tiles);
        if (s > num_kb / 16) s = num_kb / 16;
    }
    if (s > MAX_SPLITS) s = MAX_SPLITS;
    if (s > num_kb) s = num_kb;
    if (s >= 2 && tiles <= WS_COUNTERS) *splits = s;
}

static int64_t ws_bytes_for(long long tiles, int splits, int block_n) {
    if (splits <= 1) return 0;
    // the counter region has a FIXED size so that partial tiles of one launch can never land where another launch
    // (with more tiles) keeps its counters; split problems have fewer tiles than SMs, far below WS_COUNTERS
    if (tiles > WS_COUNTERS) return -1;
    return WS_COUNTERS * 4 + (int64_t)tiles * splits * BLOCK_M * block_n * 4;
}

template <int BLOCK_N>
static int32_t launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmOut,
                           const GemmKernelParams& kp, cudaStream_t stream) {
    using Cfg = GemmCfg<BLOCK_N>;
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(gemm_kernel<BLOCK_N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "gemm smem attribute: %s", cudaGetErrorString(e));
        attr_set[dev] = true;
    }
    long long total = (long long)kp.m_tiles * kp.n_tiles * kp.splits;
    int grid = (int)(total < num_sms() ? total : num_sms());
    B200SD_LAUNCH("gemm_kernel", (gemm_kernel<BLOCK_N, false>), dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, tmA, tmB,
                  tmOut, kp);
    count_launch();
    B200SD_CHECK_LAUNCH("gemm_kernel");
    return B200SD_OK;
}

// CTA-pair variant (tcgen05.mma.cta_group::2, 256 x BLOCK_N tiles): launched as 2-CTA clusters
template <int BLOCK_N>
static int32_t launch_gemm_pair(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmOut,
                                const GemmKernelParams& kp, cudaStream_t stream) {
    using Cfg = GemmCfg<BLOCK_N, true>;
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(gemm_kernel<BLOCK_N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "gemm (pair) smem attribute: %s", cudaGetErrorString(e));
        attr_set[dev] = true;
    }
    const long long pairs = (long long)((kp.m_tiles + 1) / 2) * kp.n_tiles;
    const int max_clusters = num_sms() / 2;
    const int clusters = (int)(pairs < max_clusters ? pairs : max_clusters);
    cudaError_t e = launch_kernel(gemm_kernel<BLOCK_N, true>, dim3(2 * clusters), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, 2,
                                  tmA, tmB, tmOut, kp);
    if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "gemm (pair) launch: %s", cudaGetErrorString(e));
    count_launch();
    B200SD_CHECK_LAUNCH("gemm_kernel<pair>");
    return B200SD_OK;
}

struct GemmPlan {
    GemmKernelParams kp;
    int pair;  // 1: CTA-pair kernel (block_n = 256 or 160)
    int block_n;
    int64_t ws_bytes;
};

static int32_t make_plan(const b200sd_gemm_params* p, bool have_ws, GemmPlan* plan);

// Slices per image of the column-sum table (4 per 128-row tile), or 0 when this problem's epilogue cannot produce it.
// Needs the staged epilogue on every chunk (plain column bias, n_out % 32 == 0, no GEGLU, no split K) and tiles that
// stay inside one image (NHWC: pixel boxes of one image; [M, K]: rows_per_image % 128 == 0).
static int tile_sum_parts(const b200sd_gemm_params* p, const GemmPlan& plan) {
    const GemmKernelParams& kp = plan.kp;
    if (p->geglu || kp.splits != 1 || p->bias_along_m || p->n_out % 32 != 0 || p->n_out < 32) return 0;
    if (p->row_bias && kp.bn != 1) return 0;
    const bool res_ok = !p->residual || p->residual_dtype == p->out_dtype ||
                        (p->residual_dtype == B200SD_F32 && p->out_dtype != B200SD_F32);
    if (!res_ok) return 0;
    if (p->img_h == 1 && p->n_img == 1) {
        if (p->rows_per_image <= 0 || p->rows_per_image % BLOCK_M != 0 || p->img_w % p->rows_per_image != 0) return 0;
        return (p->rows_per_image / BLOCK_M) * 4;
    }
    if (kp.bn != 1 || p->rows_per_image != 0) return 0;
    return kp.tiles_x * kp.tiles_y * 4;
}

static int32_t make_plan(const b200sd_gemm_params* p, bool have_ws, GemmPlan* plan) {
    B200SD_CHECK_ARG(p->taps == 1 || p->taps == 9, "gemm: taps must be 1 or 9 (got %d)", p->taps);
    B200SD_CHECK_ARG(p->n_img > 0 && p->img_h > 0 && p->img_w > 0 && p->cin > 0 && p->n_out > 0, "gemm: bad shape");
    B200SD_CHECK_ARG(p->cin % 8 == 0, "gemm: cin (%d) must be a multiple of 8", p->cin);
    GemmKernelParams& kp = plan->kp;
    kp = GemmKernelParams{};
    B200SD_CHECK_ARG(p->conv_stride >= 0 && p->conv_stride <= 2 && (p->conv_pad_mode == 0 || p->conv_pad_mode == 1),
                     "gemm: conv_stride %d / conv_pad_mode %d", p->conv_stride, p->conv_pad_mode);
    B200SD_CHECK_ARG((p->conv_stride <= 1 && p->conv_pad_mode == 0) || p->taps == 9,
                     "gemm: conv_stride / conv_pad_mode apply to 3x3 convolutions only");
    kp.conv_stride = p->conv_stride == 2 ? 2 : 1;
    kp.pad_lo = p->conv_pad_mode == 1 ? 0 : 1;
    int out_h = p->img_h, out_w = p->img_w;
    if (p->taps == 9 && (kp.conv_stride == 2 || kp.pad_lo == 0)) {
        B200SD_CHECK_ARG(p->img_h + kp.pad_lo >= 2 && p->img_w + kp.pad_lo >= 2, "gemm: image smaller than the filter");
        out_h = (p->img_h + kp.pad_lo + 1 - 3) / kp.conv_stride + 1;
        out_w = (p->img_w + kp.pad_lo + 1 - 3) / kp.conv_stride + 1;
    }
    kp.n_img = p->n_img, kp.img_h = out_h, kp.img_w = out_w, kp.cin = p->cin, kp.taps = p->taps;
    choose_box(p->n_img, out_h, out_w, &kp.bw, &kp.bh, &kp.bn);
    kp.tiles_x = (out_w + kp.bw - 1) / kp.bw;
    kp.tiles_y = (out_h + kp.bh - 1) / kp.bh;
    const int tiles_n = (p->n_img + kp.bn - 1) / kp.bn;
    kp.m_tiles = kp.tiles_x * kp.tiles_y * tiles_n;
    kp.kb_per_tap = (p->cin + BLOCK_K - 1) / BLOCK_K;
    kp.geglu = p->geglu ? 1 : 0;
    kp.act = p->act;
    B200SD_CHECK_ARG(p->act == 0 || (p->act == 1 && !p->geglu), "gemm: act must be 0, or 1 (GELU) without geglu");
    kp.n_out = p->n_out, kp.n_cols = kp.geglu ? p->n_out / 2 : p->n_out;
    kp.a_box_bytes = (uint32_t)(kp.bw * kp.bh * kp.bn) * BLOCK_K * 2;
    B200SD_CHECK_ARG(p->block_n == 0 || p->block_n == 64 || p->block_n == 128 || p->block_n == 160 || p->block_n == 256,
                     "gemm: block_n %d not in {0,64,128,160,256}", p->block_n);
    B200SD_CHECK_ARG(p->split_k >= 0 && p->split_k <= MAX_SPLITS, "gemm: split_k %d not in [0,%d]", p->split_k,
                     MAX_SPLITS);
    int block_n = 0, splits = 1;
    plan_tiles(p->n_out, kp.geglu, kp.m_tiles, p->taps * kp.kb_per_tap, p->block_n, p->split_k, have_ws, &block_n,
               &splits);
    if (kp.geglu) {
        B200SD_CHECK_ARG((block_n == 128 || block_n == 256) && p->n_out % block_n == 0,
                         "gemm: GEGLU needs block_n in {128,256} dividing n_out (%d)", p->n_out);
        B200SD_CHECK_ARG(!p->bias_along_m, "gemm: GEGLU with bias_along_m is not supported");
    }
    // CTA pairs: wide outputs with enough 256-row tiles to fill the machine; B200SD_GEMM_PAIR=0 switches them off
    static const char* pair_env = getenv("B200SD_GEMM_PAIR");
    const bool pair_allowed = !(pair_env && pair_env[0] == '0');
    static const long long pair_min_kb = [] { const char* e = getenv("B200SD_GEMM_PAIR_MIN_KB"); return e ? atoll(e) : 12LL; }();
    const bool wide = kp.geglu ? (p->n_out % 256 == 0) : (p->n_out >= 512 || p->n_out % 256 == 0);
    // wide linear layers whose epilogue only stores (TMA store, nothing read back) also gain from pairs at K = 320..704
    // (measured: 294912 x 960 x 320 231 -> 209 us); with a staged residual epilogue they lose (190 -> 203 us)
    // (a residual that is also the output counts: it is added by a TMA reduce, the SM reads nothing back --
    //  73728 x 640 x 640 in place 75.8 -> 67.6 us)
    const bool reduce_only = p->residual && p->residual == p->out && p->ld_res == p->ld_out &&
                             p->residual_dtype == p->out_dtype && !p->act;
    const bool store_only = wide && p->taps == 1 && p->img_h == 1 && p->n_img == 1 && (!p->residual || reduce_only) &&
                            !p->tile_sums && !p->bias_along_m;
    // 320 output channels (the 96x96 level) pair up on 160-column tiles
    const bool wide160 = !kp.geglu && !wide && p->n_out >= 320 && p->n_out % 160 == 0;
    int pair_bn = wide160 ? 160 : 256;
    const long long clusters = num_sms() / 2;
    if (wide && !kp.geglu && p->n_out % 160 == 0) {
        // both widths tile the output exactly (1280 = 5 x 256 = 8 x 160): take the one with the lower waves x per-CTA
        // operand rows, the same fetch-bound model as pick_block_n_auto.  The 12x12 level at batch 16 (36 row tiles) is
        // 90 pair tiles of 256 = 1.2 waves of 74 clusters, but 144 of 160 = 1.95 waves.
        const long long pm = (kp.m_tiles + 1) / 2;
        const long long t256 = pm * ((p->n_out + 255) / 256), t160 = pm * (p->n_out / 160);
        const long long c256 = ((t256 + clusters - 1) / clusters) * (128 + 128), c160 = ((t160 + clusters - 1) / clusters) * (128 + 80);
        if (c160 < c256) pair_bn = 160;
    }
    const long long pair_tiles = (long long)((kp.m_tiles + 1) / 2) * ((p->n_out + pair_bn - 1) / pair_bn);
    // ... and long enough to pay for the cluster start-up: at least one full wave of clusters and >= 64 k-blocks per
    // cluster in total (measured at batch 1: launches of 20-50 us lost 1-5 us each as pairs, the long ones gained 10-18 %)
    const long long num_kb_total = (long long)p->taps * kp.kb_per_tap;
    plan->pair = (pair_allowed && (wide || wide160) && splits == 1 && (p->block_n == 0 || p->block_n == pair_bn) &&
                  kp.m_tiles >= 2 &&
                  pair_tiles >= clusters && pair_tiles * num_kb_total >= 64 * clusters &&
                  (num_kb_total >= pair_min_kb || kp.geglu || (store_only && num_kb_total >= 5))) ? 1 : 0;  // K < 768 with a residual epilogue: measured 10-20 % slower
                                                              // as pairs (the epilogue, not the operand fetch, is the bound)
    if (plan->pair) block_n = pair_bn;
    kp.n_tiles = (p->n_out + block_n - 1) / block_n;
    kp.splits = splits;
    plan->block_n = block_n;
    plan->ws_bytes = ws_bytes_for((long long)kp.m_tiles * kp.n_tiles, splits, block_n);
    return B200SD_OK;
}

}  // namespace b200sd

using namespace b200sd;

extern "C" int32_t b200sd_gemm_pick_block_n(int32_t n_out, int32_t geglu) {
    return pick_block_n_auto(n_out, geglu, 1 << 20);
}

extern "C" int32_t b200sd_gemm_tile_sum_parts(const b200sd_gemm_params* p) {
    if (!p) return 0;
    GemmPlan plan;
    {
        // an under-filled problem is better served by splitting K (which rules the column sums out: they come from
        // whole-K accumulators) -- its output is small, so the statistics pass it then needs is cheap
        b200sd_gemm_params w = *p;
        w.split_k = 0;
        if (make_plan(&w, true, &plan) == B200SD_OK && plan.kp.splits > 1) return 0;
    }
    // planned as b200sd_gemm will plan it when it is handed tile_sums: never split along K
    b200sd_gemm_params q = *p;
    q.split_k = 1;
    if (make_plan(&q, false, &plan) != B200SD_OK) return 0;
    return tile_sum_parts(&q, plan);
}

extern "C" int64_t b200sd_gemm_workspace_bytes(const b200sd_gemm_params* p) {
    if (!p) return 0;
    GemmPlan plan;
    if (make_plan(p, true, &plan) != B200SD_OK) return 0;
    return plan.ws_bytes;
}

extern "C" int32_t b200sd_gemm(const b200sd_gemm_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->a && p->w && p->out, "gemm: null pointer");
    B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(p->a) & 15) == 0 && (reinterpret_cast<uintptr_t>(p->w) & 15) == 0,
                     "gemm: a / w must be 16-byte aligned");
    B200SD_CHECK_ARG(p->out_dtype >= B200SD_BF16 && p->out_dtype <= B200SD_F16, "gemm: bad out_dtype");
    B200SD_CHECK_ARG(!p->residual || (p->residual_dtype >= B200SD_BF16 && p->residual_dtype <= B200SD_F16),
                     "gemm: bad residual_dtype");
    B200SD_CHECK_ARG((p->a_dtype == B200SD_BF16 || p->a_dtype == B200SD_F16) &&
                         (p->w_dtype == B200SD_BF16 || p->w_dtype == B200SD_F16),
                     "gemm: a_dtype / w_dtype must be bf16 or fp16");
    // measured on B200: tcgen05.mma kind::f16 with A and B in different 16-bit formats traps (illegal instruction)
    if (p->a_dtype != p->w_dtype)
        return set_error(B200SD_ERR_UNSUPPORTED, "gemm: a_dtype (%d) and w_dtype (%d) must be equal", p->a_dtype,
                         p->w_dtype);
    const int n_cols = p->geglu ? p->n_out / 2 : p->n_out;
    if (n_cols >= 32) {
        // vector path requirements
        const int oa = p->out_dtype == B200SD_F32 ? 4 : 8;
        B200SD_CHECK_ARG(p->ld_out % oa == 0 && (reinterpret_cast<uintptr_t>(p->out) & 15) == 0,
                         "gemm: out must be 16-byte aligned with ld_out %% %d == 0", oa);
        if (p->residual) {
            const int ra = p->residual_dtype == B200SD_F32 ? 4 : 8;
            B200SD_CHECK_ARG(p->ld_res % ra == 0 && (reinterpret_cast<uintptr_t>(p->residual) & 15) == 0,
                             "gemm: residual must be 16-byte aligned with ld_res %% %d == 0", ra);
        }
        if (p->bias && !p->bias_along_m)
            B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(p->bias) & 15) == 0, "gemm: bias must be 16-byte aligned");
        if (p->row_bias)
            B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(p->row_bias) & 15) == 0 && p->row_bias_stride % 4 == 0,
                             "gemm: row_bias must be 16-byte aligned, stride %% 4 == 0");
    }
    const bool have_ws = p->workspace != nullptr && p->workspace_bytes > 0;
    B200SD_CHECK_ARG(!have_ws || (reinterpret_cast<uintptr_t>(p->workspace) & 255) == 0,
                     "gemm: workspace must be 256-byte aligned");
    GemmPlan plan;
    int32_t rc = make_plan(p, have_ws && !p->tile_sums, &plan);  // column sums come from whole-K accumulators
    if (rc) return rc;
    if (p->tile_sums) {
        B200SD_CHECK_ARG(p->split_k <= 1, "gemm: tile_sums cannot be combined with split_k");
        B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(p->tile_sums) & 15) == 0, "gemm: tile_sums must be 16-byte aligned");
        if (tile_sum_parts(p, plan) == 0)
            return set_error(B200SD_ERR_UNSUPPORTED, "gemm: this problem cannot emit tile_sums (ask "
                             "b200sd_gemm_tile_sum_parts first): tiles span images, or the epilogue is not the staged one");
    }
    if (plan.kp.splits > 1 && plan.ws_bytes > p->workspace_bytes) {
        if (p->split_k > 1)
            return set_error(B200SD_ERR_WORKSPACE, "gemm: split_k=%d needs %lld workspace bytes, have %lld", p->split_k,
                             (long long)plan.ws_bytes, (long long)p->workspace_bytes);
        rc = make_plan(p, false, &plan);  // workspace too small for the automatic split: run unsplit
        if (rc) return rc;
    }
    GemmKernelParams& kp = plan.kp;
    const int block_n = plan.block_n;
    kp.a_bf16 = p->a_dtype == B200SD_BF16, kp.b_bf16 = p->w_dtype == B200SD_BF16;
    kp.bias = p->bias, kp.bias_along_m = p->bias_along_m;
    kp.row_bias = p->row_bias, kp.row_bias_stride = p->row_bias_stride;
    kp.residual = p->residual, kp.res_dtype = p->residual_dtype, kp.ld_res = p->ld_res;
    kp.out = p->out, kp.out_dtype = p->out_dtype, kp.ld_out = p->ld_out;
    {
        // staged (transposed) epilogue: column bias only, residual of the output's dtype; 16-byte alignment
        // of out / residual rows was checked above for n_cols >= 32
        const bool res_ok = !p->residual || p->residual_dtype == p->out_dtype ||
                            (p->residual_dtype == B200SD_F32 && p->out_dtype != B200SD_F32);
        // (a row bias is folded into the tile's column bias when a tile never spans two images)
        kp.fast_epi = (n_cols >= 32 && !p->bias_along_m && (!p->row_bias || kp.bn == 1) && res_ok) ? 1 : 0;
        // TMA-store epilogue: plain [M, n_cols] row-major output (linear layers), no residual, unsplit
        static const bool tma_off = getenv("B200SD_GEMM_NO_TMA_EPI") != nullptr;
        const int osz = p->out_dtype == B200SD_F32 ? 4 : 2;
        // An fp32 residual that IS the output (h += linear(x), the transformer's projections back onto the trunk) needs no
        // residual read at all: the chunk (acc + bias) is added to it in L2 by a TMA reduce -- same fp32 sum, same order.
        static const bool red_off = getenv("B200SD_GEMM_NO_TMA_REDUCE") != nullptr;
        // The same holds for a 16-bit trunk (the default): the chunk is rounded to the trunk's format and added to it in
        // L2 -- two roundings (chunk, then sum) where the read path has one; B200SD_GEMM_NO_TMA_REDUCE16 keeps the read path.
        static const bool red16_off = getenv("B200SD_GEMM_NO_TMA_REDUCE16") != nullptr;
        const bool inplace32 = p->residual == p->out && p->ld_res == p->ld_out && p->residual_dtype == p->out_dtype &&
                               (p->out_dtype == B200SD_F32 || !red16_off) && !p->act && !red_off;
        // (column sums are taken in the staged epilogue's transposed read-back, so they rule the TMA store out)
        kp.tma_epi = (kp.fast_epi && !tma_off && (!p->residual || inplace32) && kp.splits == 1 && p->taps == 1 &&
                      p->img_h == 1 && p->n_img == 1 && (p->ld_out * osz) % 16 == 0 && !p->tile_sums) ? 1 : 0;
        kp.tile_sums = reinterpret_cast<float2*>(p->tile_sums);
        kp.tma_reduce = (kp.tma_epi && p->residual) ? 1 : 0;
        static const bool res_pf = [] { const char* e = getenv("B200SD_GEMM_RES_PREFETCH"); return e && e[0] == '1'; }();
        kp.res_prefetch = res_pf ? 1 : 0;
    }
    if (kp.splits > 1) {
        const long long tiles = (long long)kp.m_tiles * kp.n_tiles;
        kp.counters = static_cast<int*>(p->workspace);
        (void)tiles;
        kp.partials = reinterpret_cast<float*>(static_cast<uint8_t*>(p->workspace) + WS_COUNTERS * 4);
    }

    CUtensorMap tmA, tmB;
    {
        const uint64_t c = (uint64_t)p->cin;
        uint64_t dims[4] = {c, (uint64_t)p->img_w, (uint64_t)p->img_h, (uint64_t)p->n_img};
        uint64_t strides[3] = {c * 2, c * 2 * p->img_w, c * 2 * p->img_w * p->img_h};
        // stride 2: the box spans 2 * bw x 2 * bh input pixels and the element strides keep every second one, so the
        // shared-memory tile is the same dense [bn x bh x bw pixels] x 64 channels as for stride 1
        const uint32_t cs = (uint32_t)kp.conv_stride;
        uint32_t box[4] = {BLOCK_K, (uint32_t)kp.bw * cs, (uint32_t)kp.bh * cs, (uint32_t)kp.bn};
        uint32_t estr[4] = {1, cs, cs, 1};
        int32_t r = encode_tmap_bf16(&tmA, p->a, 4, dims, strides, box, estr);
        if (r) return r;
    }
    {
        const uint64_t k_total = (uint64_t)p->taps * p->cin;
        uint64_t dims[2] = {k_total, (uint64_t)p->n_out};
        uint64_t strides[1] = {k_total * 2};
        uint32_t box[2] = {BLOCK_K, (uint32_t)(plan.pair ? block_n / 2 : block_n)};  // a pair's CTA stages half of the tile's rows
        int32_t r = encode_tmap_bf16(&tmB, p->w, 2, dims, strides, box);
        if (r) return r;
    }
    CUtensorMap tmOut = tmB;  // placeholder when the TMA-store epilogue is off (never dereferenced then)
    if (kp.tma_epi) {
        const int osz = p->out_dtype == B200SD_F32 ? 4 : 2;
        int32_t r = encode_tmap_2d(&tmOut, p->out, osz, (uint64_t)n_cols, (uint64_t)p->img_w,
                                   (uint64_t)p->ld_out * osz, 32, 32, osz == 4 ? 128 : 64, p->out_dtype);
        if (r) return r;
    }
    cudaStream_t s = as_stream(stream);
    if (plan.pair)
        return block_n == 160 ? launch_gemm_pair<160>(tmA, tmB, tmOut, kp, s) : launch_gemm_pair<256>(tmA, tmB, tmOut, kp, s);
    switch (block_n) {
        case 64: return launch_gemm<64>(tmA, tmB, tmOut, kp, s);
        case 128: return launch_gemm<128>(tmA, tmB, tmOut, kp, s);
        case 160: return launch_gemm<160>(tmA, tmB, tmOut, kp, s);
        default: return launch_gemm<256>(tmA, tmB, tmOut, kp, s);
    }
}

// ------------------------------------------------------------------------------------------ UMMA self-test
namespace b200sd {

// Single CTA, 128 threads.  Operands are written to shared memory by ordinary stores in the SWIZZLE_128B
// pattern (16-byte chunk c of 128-byte row r lives at chunk c ^ (r & 7)), then consumed by tcgen05.mma.
__global__ void __launch_bounds__(128, 1)
    umma_selftest_kernel(const uint16_t* __restrict__ A, const uint16_t* __restrict__ B, float* __restrict__ D,
                         int n, int k, int b_mn_major, int a_bf16, int b_bf16) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
    uint8_t* sA = smem;                  // k/64 tiles of [128 rows x 128 B]
    uint8_t* sB = smem + 128 * k * 2;    // K-major: k/64 tiles of [n rows x 128 B]; MN-major: n/64 atoms of [k rows x 128 B]
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // A: row r, 16-byte chunk index cc (8 elements) over the whole K extent
    for (int i = tid; i < 128 * (k / 8); i += 128) {
        int r = i / (k / 8), cc = i % (k / 8);
        int kc = cc / 8, c = cc % 8;
        uint4 v = *reinterpret_cast<const uint4*>(A + (size_t)r * k + cc * 8);
        *reinterpret_cast<uint4*>(sA + kc * 16384 + r * 128 + ((c ^ (r & 7)) << 4)) = v;
    }
    if (!b_mn_major) {
        for (int i = tid; i < n * (k / 8); i += 128) {
            int r = i / (k / 8), cc = i % (k / 8);
            int kc = cc / 8, c = cc % 8;
            uint4 v = *reinterpret_cast<const uint4*>(B + (size_t)r * k + cc * 8);
            *reinterpret_cast<uint4*>(sB + kc * (n * 128) + r * 128 + ((c ^ (r & 7)) << 4)) = v;
        }
    } else {
        // B given as [k, n] (n contiguous): atom na covers n-columns [64 na, 64 na + 64); row = k index
        for (int i = tid; i < k * (n / 8); i += 128) {
            int r = i / (n / 8), cc = i % (n / 8);
            int na = cc / 8, c = cc % 8;
            uint4 v = *reinterpret_cast<const uint4*>(B + (size_t)r * n + cc * 8);
            *reinterpret_cast<uint4*>(sB + na * (k * 128) + r * 128 + ((c ^ (r & 7)) << 4)) = v;
        }
    }
    fence_proxy_async_smem();
    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (tid == 0) {
        const uint32_t idesc = umma_idesc(128, (uint32_t)n, (uint32_t)a_bf16, (uint32_t)b_bf16, 0, b_mn_major ? 1u : 0u);
        for (int kk = 0; kk < k / 16; ++kk) {
            const uint32_t a_addr = smem_u32(sA) + (kk / 4) * 16384 + (kk % 4) * 32;
            uint64_t adesc = umma_desc_sw128(a_addr);
            uint64_t bdesc;
            if (!b_mn_major)
                bdesc = umma_desc_sw128(smem_u32(sB) + (kk / 4) * (n * 128) + (kk % 4) * 32);
            else
                bdesc = umma_desc_sw128(smem_u32(sB) + kk * 2048, /*lbo=*/(uint32_t)k * 128, /*sbo=*/1024);
            umma_bf16(tmem_base, adesc, bdesc, idesc, kk != 0);
        }
        umma_commit(&bar);
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    tc_fence_after();
    for (int c = 0; c < n / 16; ++c) {
        uint32_t r[16];
        tmem_ld_x16(tmem_base + c * 16 + ((uint32_t)(warp * 32) << 16), r);
        tmem_ld_wait();
        for (int j = 0; j < 16; ++j) D[(size_t)(warp * 32 + lane) * n + c * 16 + j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

}  // namespace b200sd

extern "C" int32_t b200sd_umma_selftest(const void* a, int32_t a_dtype, const void* b, int32_t b_dtype, float* d,
                                        int32_t n, int32_t k, int32_t b_mn_major, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(a && b && d, "umma_selftest: null pointer");
    B200SD_CHECK_ARG((a_dtype == B200SD_BF16 || a_dtype == B200SD_F16) && (b_dtype == B200SD_BF16 || b_dtype == B200SD_F16),
                     "umma_selftest: operands must be bf16 or fp16");
    if (a_dtype != b_dtype)
        return set_error(B200SD_ERR_UNSUPPORTED, "umma_selftest: mixed bf16 x fp16 operands trap on sm_100a");
    B200SD_CHECK_ARG(n >= 16 && n <= 256 && n % 16 == 0 && k >= 64 && k <= 256 && k % 64 == 0,
                     "umma_selftest: need 16<=n<=256 (n%%16==0), 64<=k<=256 (k%%64==0)");
    B200SD_CHECK_ARG(!b_mn_major || n % 64 == 0, "umma_selftest: MN-major B needs n %% 64 == 0");
    const int smem = 128 * k * 2 + n * k * 2 + 1024;
    cudaError_t e =
        cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "umma_selftest attr: %s", cudaGetErrorString(e));
    umma_selftest_kernel<<<1, 128, smem, as_stream(stream)>>>(static_cast<const uint16_t*>(a),
                                                              static_cast<const uint16_t*>(b), d, n, k, b_mn_major,
                                                              a_dtype == B200SD_BF16, b_dtype == B200SD_BF16);
    count_launch();
    B200SD_CHECK_LAUNCH("umma_selftest_kernel");
    return B200SD_OK;
}
