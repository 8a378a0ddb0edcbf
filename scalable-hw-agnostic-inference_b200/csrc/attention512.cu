// Flash attention for ONE head of dimension 512 on sm_100a: the mid-block attention of AutoencoderKL (decoder and
// encoder), which diffusers runs as softmax(Q K^T / sqrt(512)) V over all h*w tokens of the latent (9216 at 768^2) and
// the round-1 path ran as GEMM -> row softmax -> GEMM with the S x S score matrix in HBM (340 MB per image).
//
// head_dim 512 does not fit the d = 64 kernel's plan (attention.cu): a 128 x 512 fp32 O accumulator is all 512 TMEM
// columns by itself, and a 128 x 512 Q tile plus a 128 x 512 K tile are 256 KiB of shared memory.  So:
//   * d is split in two for O: CTA (q tile, half) computes the whole score tile S = Q K^T (all 512 dims) but only 256 of
//     the 512 output columns (O += P V[:, half]).  TMEM: S 128 columns, P 64, O 256.  The score work is done twice
//     (1.5x the FLOPs of one pass); this layer is 1.5 % of a VAE decode.
//   * the d = 512 contraction of S streams through shared memory in eight 64-wide chunks (Q_c, K_c: 16 KiB each, a
//     2-stage ring); Q is re-streamed from L2 for every kv tile instead of being held.
//   * V's half tile (128 kv rows x 256 dims = 64 KiB, four MN-major atoms) has its own 2-stage ring.
// Warp roles as in attention.cu: warp 0 TMA, warp 1 UMMA (in order: Q K^T of step j+1, then P V of step j), warps 2..9
// softmax with two threads per query row; P stays in TMEM and feeds the P V UMMA as its A operand.
// FLOPs per launch (algorithmic, one pass): 4 * batch * sq * skv * 512.
#include "common.h"
#include "ptx.cuh"

namespace b200sd {
using namespace ptx;

constexpr int A5_BM = 128;             // query rows per CTA
constexpr int A5_BN = 128;             // kv rows per tile
constexpr int A5_D = 512;
constexpr int A5_DH = 256;             // output columns per CTA
constexpr int A5_CHUNKS = A5_D / 64;   // 64-wide chunks of the Q K^T contraction
constexpr int A5_TILE = 128 * 128;     // bytes of a [128 x 64] 16-bit tile
constexpr int A5_QK_STAGES = 2;        // (Q chunk, K chunk) pairs
constexpr int A5_V_STAGES = 2;         // V half tiles (4 tiles of A5_TILE)
constexpr int A5_THREADS = 32 * 10;
constexpr int A5_XCH = 3 * 256 * 4;    // row-max exchange [parity][half][128] + row-sum exchange [half][128], fp32
constexpr int A5_SMEM = A5_QK_STAGES * 2 * A5_TILE + A5_V_STAGES * 4 * A5_TILE + A5_XCH + 1024 + 256;
constexpr uint32_t A5_TMEM_S = 0, A5_TMEM_P = 128, A5_TMEM_O = 256;

struct Attn512Params {
    int sq, skv, n_kv_tiles;
    float scale_log2;
    uint16_t* out;
    long long ld_out;
};

__device__ __forceinline__ float a5_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float a5_max3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void a5_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void a5_tmem_st_x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}

template <bool F16>
__global__ void __launch_bounds__(A5_THREADS, 1)
    attention512_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const Attn512Params p) {
    constexpr float TAU = 8.0f;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
    uint8_t* sQK = smem;                                      // stage s: Q chunk at s*2*TILE, K chunk at s*2*TILE + TILE
    uint8_t* sV = sQK + A5_QK_STAGES * 2 * A5_TILE;           // stage s: four [128 kv x 64 d] atoms
    float* xchg = reinterpret_cast<float*>(sV + A5_V_STAGES * 4 * A5_TILE);
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(xchg) + A5_XCH);
    uint64_t* qk_full = bars;
    uint64_t* qk_empty = qk_full + A5_QK_STAGES;
    uint64_t* v_full = qk_empty + A5_QK_STAGES;
    uint64_t* v_empty = v_full + A5_V_STAGES;
    uint64_t* s_full = v_empty + A5_V_STAGES;   // UMMA -> softmax: S tile complete
    uint64_t* s_free = s_full + 1;              // softmax -> UMMA: S tile copied to registers
    uint64_t* p_full = s_free + 1;              // softmax -> UMMA: P tile in TMEM (and O rescaled)
    uint64_t* pv_done = p_full + 1;             // UMMA -> softmax: O += P V finished
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = p.n_kv_tiles;
    const int q0 = blockIdx.x * A5_BM, dh = blockIdx.y, b = blockIdx.z;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < A5_QK_STAGES; ++s) mbar_init(&qk_full[s], 1), mbar_init(&qk_empty[s], 1);
        for (int s = 0; s < A5_V_STAGES; ++s) mbar_init(&v_full[s], 1), mbar_init(&v_empty[s], 1);
        mbar_init(s_full, 1);
        mbar_init(s_free, 8);
        mbar_init(p_full, 8);
        mbar_init(pv_done, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_sync();

    constexpr uint32_t FULL = 0xffffffffu;
    if (warp == 0) {
        // ---------------------------------------------------------------- TMA producer
        int qs = 0, vs = 0;
        uint32_t qph = 0, vph = 0;
        for (int j = 0; j < n_tiles; ++j) {
            for (int c = 0; c < A5_CHUNKS; ++c) {
                mbar_wait(&qk_empty[qs], qph ^ 1);
                uint8_t* dst = sQK + qs * 2 * A5_TILE;
                if (elect_one()) {
                    mbar_expect_tx(&qk_full[qs], 2 * A5_TILE);
                    tma_load_2d(dst, &tmQ, &qk_full[qs], c * 64, b * p.sq + q0);
                    tma_load_2d(dst + A5_TILE, &tmK, &qk_full[qs], c * 64, b * p.skv + j * A5_BN);
                }
                __syncwarp();
                if (++qs == A5_QK_STAGES) qs = 0, qph ^= 1;
            }
            mbar_wait(&v_empty[vs], vph ^ 1);
            uint8_t* dv = sV + vs * 4 * A5_TILE;
            if (elect_one()) {
                mbar_expect_tx(&v_full[vs], 4 * A5_TILE);
#pragma unroll
                for (int a = 0; a < 4; ++a)
                    tma_load_2d(dv + a * A5_TILE, &tmV, &v_full[vs], dh * A5_DH + a * 64, b * p.skv + j * A5_BN);
            }
            __syncwarp();
            if (++vs == A5_V_STAGES) vs = 0, vph ^= 1;
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- UMMA issue, in order
        constexpr uint32_t bf = F16 ? 0u : 1u;
        constexpr uint32_t idesc_qk = umma_idesc(128, A5_BN, bf, bf, 0, 0);
        constexpr uint32_t idesc_pv = umma_idesc(128, A5_DH, bf, bf, 0, 1);  // A = P (TMEM), B = V MN-major, N = 256
        const uint32_t tmem_u = __shfl_sync(FULL, tmem_base, 0);
        const uint32_t qk_u = __shfl_sync(FULL, smem_u32(sQK), 0);
        const uint32_t v_u = __shfl_sync(FULL, smem_u32(sV), 0);
        const uint32_t s_tmem = tmem_u + A5_TMEM_S, p_tmem = tmem_u + A5_TMEM_P, o_tmem = tmem_u + A5_TMEM_O;
        int qs = 0, vs = 0;
        uint32_t qph = 0, vph = 0;
        auto issue_qk = [&](int j) {
            if (j > 0) mbar_wait(s_free, (j - 1) & 1);   // the previous S tile sits in the softmax warps' registers
            for (int c = 0; c < A5_CHUNKS; ++c) {
                mbar_wait(&qk_full[qs], qph);
                tc_fence_after();
                const uint64_t qdesc = umma_desc_sw128(qk_u + qs * 2 * A5_TILE);
                const uint64_t kdesc = umma_desc_sw128(qk_u + qs * 2 * A5_TILE + A5_TILE);
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_bf16(s_tmem, qdesc + 2 * ks, kdesc + 2 * ks, idesc_qk, (c | ks) != 0);
                    umma_commit(&qk_empty[qs]);
                    if (c == A5_CHUNKS - 1) umma_commit(s_full);
                }
                __syncwarp();
                if (++qs == A5_QK_STAGES) qs = 0, qph ^= 1;
            }
        };
        auto issue_pv = [&](int j) {
            mbar_wait(p_full, j & 1);
            mbar_wait(&v_full[vs], vph);
            tc_fence_after();
            // V half tile: four MN-major atoms of [128 kv x 64 d] 16 KiB apart (LBO), 8-row groups 1 KiB apart (SBO);
            // 16 kv rows per UMMA = 2 KiB further down each atom
            const uint32_t v_addr = v_u + vs * 4 * A5_TILE;
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < A5_BN / 16; ++ks)
                    a5_umma_ts(o_tmem, p_tmem + ks * 8, umma_desc_sw128(v_addr + ks * 2048, /*lbo*/ A5_TILE, /*sbo*/ 1024),
                               idesc_pv, (j | ks) != 0);
                umma_commit(pv_done);
                umma_commit(&v_empty[vs]);
            }
            __syncwarp();
            if (++vs == A5_V_STAGES) vs = 0, vph ^= 1;
        };
        issue_qk(0);
        for (int j = 0; j < n_tiles; ++j) {
            if (j + 1 < n_tiles) issue_qk(j + 1);
            issue_pv(j);
        }
    } else {
        // ---------------------------------------------------------------- softmax + output (8 warps)
        // warps 2..5 own score columns 0..63 of their 32 rows, warps 6..9 columns 64..127 of the same rows
        const int half = ((warp - 2) >> 2) & 1;
        const int quad = warp & 3;
        const int row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t s_addr = tmem_base + A5_TMEM_S + half * 64 + lane_off;
        const uint32_t p_addr = tmem_base + A5_TMEM_P + half * 32 + lane_off;
        const uint32_t o_addr = tmem_base + A5_TMEM_O + half * 128 + lane_off;   // this thread's 128 of the 256 O columns
        const float sc = p.scale_log2;
        float m_used = 0.f, l = 0.f;
        for (int j = 0; j < n_tiles; ++j) {
            mbar_wait(s_full, j & 1);
            tc_fence_after();
            uint32_t s[2][32];
#pragma unroll
            for (int c = 0; c < 2; ++c) tmem_ld_x32(s_addr + c * 32, s[c]);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_free);
            const int valid = p.skv - j * A5_BN - half * 64;   // live kv columns among this thread's 64
            if (valid < 64) {                                  // ragged last kv tile: warp-uniform
                __syncwarp();
#pragma unroll
                for (int c = 0; c < 2; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c * 32 + i >= valid) s[c][i] = __float_as_uint(-INFINITY);
            }
            float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    mx0 = a5_max3(mx0, __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
                    mx1 = a5_max3(mx1, __uint_as_float(s[c][i + 2]), __uint_as_float(s[c][i + 3]));
                }
            float* xch = xchg + (j & 1) * 256;
            xch[half * 128 + row] = fmaxf(mx0, mx1);
            named_bar_sync(4 + quad, 64);   // the two warps that share these 32 rows
            const float m_new = fmaxf(xch[row], xch[128 + row]) * sc;
            float alpha = 1.f;
            bool need = false;
            if (j == 0) {
                m_used = m_new;
            } else if (m_new > m_used + TAU) {
                alpha = a5_ex2(m_used - m_new);
                m_used = m_new;
                need = true;
            }
            uint32_t pk[32];
            const uint64_t sc2 = pack2(sc, sc), negm2 = pack2(-m_used, -m_used);
            uint64_t ls2 = pack2(0.f, 0.f);
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const uint64_t x2 = fma2(pack2(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1])), sc2, negm2);
                    float x0, x1;
                    unpack2(x2, x0, x1);
                    const float e0 = a5_ex2(x0), e1 = a5_ex2(x1);
                    ls2 = add2(ls2, pack2(e0, e1));
                    pk[c * 16 + i] = pack_h2(e0, e1, F16);
                }
            float ls0, ls1;
            unpack2(ls2, ls0, ls1);
            if (j > 0) {   // P V of the previous step must be complete before P is overwritten or O rescaled
                mbar_wait(pv_done, (j - 1) & 1);
                tc_fence_after();
            }
            if (__any_sync(FULL, need)) {
#pragma unroll 1
                for (int c = 0; c < 8; ++c) {   // this thread's 128 O columns, 16 at a time
                    uint32_t o[16];
                    tmem_ld_x16(o_addr + c * 16, o);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    tmem_st_x16(o_addr + c * 16, o);
                }
            }
            l = l * alpha + (ls0 + ls1);
            a5_tmem_st_x32(p_addr, pk);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        // ---- epilogue: O / l -> 16-bit
        float* xch = xchg + 2 * 256;
        xch[half * 128 + row] = l;
        named_bar_sync(4 + quad, 64);
        const float inv = 1.0f / (xch[row] + xch[128 + row]);
        mbar_wait(pv_done, (n_tiles - 1) & 1);
        tc_fence_after();
        const int qrow = q0 + row;
        uint16_t* orow = p.out + ((long long)b * p.sq + qrow) * p.ld_out + dh * A5_DH + half * 128;
#pragma unroll 1
        for (int c = 0; c < 8; ++c) {
            uint32_t o[16];
            tmem_ld_x16(o_addr + c * 16, o);
            tmem_ld_wait();
            if (qrow < p.sq) {
                uint32_t w[8];
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    w[i] = pack_h2(__uint_as_float(o[2 * i]) * inv, __uint_as_float(o[2 * i + 1]) * inv, F16);
                *reinterpret_cast<uint4*>(orow + c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
                *reinterpret_cast<uint4*>(orow + c * 16 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

template <bool F16>
static int32_t launch_attention512(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                   const Attn512Params& kp, dim3 grid, cudaStream_t stream) {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(attention512_kernel<F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, A5_SMEM);
        if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "attention512 smem attribute: %s", cudaGetErrorString(e));
        attr_set[dev] = true;
    }
    B200SD_LAUNCH("attention512_kernel", (attention512_kernel<F16>), grid, dim3(A5_THREADS), A5_SMEM, stream, tmQ, tmK, tmV, kp);
    count_launch();
    B200SD_CHECK_LAUNCH("attention512_kernel");
    return B200SD_OK;
}

}  // namespace b200sd

using namespace b200sd;

extern "C" int32_t b200sd_attention512(const b200sd_attention_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->q && p->k && p->v && p->out, "attention512: null pointer");
    if (p->head_dim != 512 || p->heads != 1)
        return set_error(B200SD_ERR_UNSUPPORTED, "attention512: one head of dimension 512 (got %d x %d)", p->heads, p->head_dim);
    B200SD_CHECK_ARG(p->batch > 0 && p->sq > 0 && p->skv > 0 && !p->causal, "attention512: bad shape / causal not supported");
    B200SD_CHECK_ARG(p->dtype == B200SD_BF16 || p->dtype == B200SD_F16, "attention512: dtype must be bf16 or fp16");
    B200SD_CHECK_ARG(p->scale > 0.f, "attention512: scale must be positive");
    B200SD_CHECK_ARG(p->ld_q % 8 == 0 && p->ld_k % 8 == 0 && p->ld_v % 8 == 0 && p->ld_out % 8 == 0,
                     "attention512: leading dimensions must be multiples of 8");
    B200SD_CHECK_ARG(((reinterpret_cast<uintptr_t>(p->q) | reinterpret_cast<uintptr_t>(p->k) |
                       reinterpret_cast<uintptr_t>(p->v) | reinterpret_cast<uintptr_t>(p->out)) & 15) == 0,
                     "attention512: pointers must be 16-byte aligned");
    B200SD_CHECK_ARG(p->batch <= 65535, "attention512: batch too large for the grid");
    CUtensorMap tmQ, tmK, tmV;
    const uint32_t box[2] = {64, 128};
    {
        uint64_t dims[2] = {(uint64_t)A5_D, (uint64_t)p->batch * p->sq};
        uint64_t str[1] = {(uint64_t)p->ld_q * 2};
        int32_t r = encode_tmap_bf16(&tmQ, p->q, 2, dims, str, box);
        if (r) return r;
    }
    {
        uint64_t dims[2] = {(uint64_t)A5_D, (uint64_t)p->batch * p->skv};
        uint64_t strk[1] = {(uint64_t)p->ld_k * 2};
        uint64_t strv[1] = {(uint64_t)p->ld_v * 2};
        int32_t r = encode_tmap_bf16(&tmK, p->k, 2, dims, strk, box);
        if (r) return r;
        r = encode_tmap_bf16(&tmV, p->v, 2, dims, strv, box);
        if (r) return r;
    }
    Attn512Params kp;
    kp.sq = p->sq, kp.skv = p->skv;
    kp.n_kv_tiles = (p->skv + A5_BN - 1) / A5_BN;
    kp.scale_log2 = p->scale * 1.4426950408889634f;
    kp.out = static_cast<uint16_t*>(p->out);
    kp.ld_out = p->ld_out;
    dim3 grid((unsigned)((p->sq + A5_BM - 1) / A5_BM), 2, (unsigned)p->batch);
    cudaStream_t s = as_stream(stream);
    return p->dtype == B200SD_F16 ? launch_attention512<true>(tmQ, tmK, tmV, kp, grid, s)
                                  : launch_attention512<false>(tmQ, tmK, tmV, kp, grid, s);
}
