// K3: fused flash attention for head_dim 64 on sm_100a (self- and cross-attention of BasicTransformerBlock).
//
// One CTA owns 256 query rows = two 128-row tiles (A, B) that share every K/V tile:
//   warp 0        TMA producer : both Q tiles once, then K/V tiles (128 kv rows x 64) through a 3-stage ring
//   warp 1        UMMA issuer  : S_t = Q_t K^T (128x128x64) into TMEM; O_t += P_t V (128x64x128) with P_t from shared
//                                memory and V as an MN-major operand (no transpose of V anywhere); L_t += P_t 1
//                                (128x16x128 against a constant tile of ones) = the softmax denominators, accumulated
//                                in fp32 by the tensor core instead of one FADD per element in the softmax warps
//   warps 2..5    softmax A    } one query row per thread (tcgen05.ld 32x32b => no cross-lane reductions), the whole
//   warps 6..9    softmax B    } score row held in registers so S_t is released to the next QK^T at once; lazy rescale
//                                of O and L (only when the running max grows by more than 2^TAU); P written as 16-bit
//                                into the SWIZZLE_128B layout UMMA reads.
// The exponentials are the bound at d = 64 (16 MUFU/clk/SM against 8192 MACs/clk/SM).  With fp16 operands the
// exponent is computed two at a time by ex2.approx.ftz.f16x2, whose packed result IS the fp16 P tile: half the MUFU
// work, no conversion.  With bf16 operands the fp32 ex2 path is kept (a bf16 exponent argument is too coarse).
// The S x S score matrix never leaves the SM.  FLOPs per launch: 4 * batch * heads * sq * skv * 64.
#include "common.h"
#include "ptx.cuh"

namespace b200sd {
using namespace ptx;

constexpr int ATT_BM = 128;          // rows per Q tile
constexpr int ATT_QT = 2;            // Q tiles per CTA
constexpr int ATT_BN = 128;          // kv rows per tile
constexpr int ATT_D = 64;
constexpr int ATT_STAGES = 3;
constexpr int ATT_TILE = 128 * 128;  // bytes of a [128 x 64] 16-bit tile
constexpr int ATT_ONES = 2 * 16 * 128;  // [16 rows x 128 k] of 1.0, two 64-wide k chunks
constexpr int ATT_THREADS = 32 * (2 + 4 * ATT_QT);
constexpr int ATT_SMEM = ATT_QT * ATT_TILE /*Q*/ + ATT_STAGES * 2 * ATT_TILE /*K,V*/ + ATT_QT * 2 * ATT_TILE /*P*/ +
                         ATT_ONES + 1024 + 256;
// TMEM columns: S_A 0..127, S_B 128..255, O_A 256..319, O_B 320..383, L_A 384..399, L_B 400..415
constexpr uint32_t ATT_TMEM_S = 0, ATT_TMEM_O = 256, ATT_TMEM_L = 384;

struct AttnKernelParams {
    int sq, skv, n_kv_tiles;
    float scale_log2;  // scale * log2(e)
    uint16_t* out;
    long long ld_out;
};

__device__ __forceinline__ float ex2_ftz(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// two exponentials at once on fp16 inputs; the packed fp16 result goes straight into the P tile
__device__ __forceinline__ uint32_t ex2_f16x2(float lo, float hi) {
    uint32_t h, y;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(hi), "f"(lo));
    asm("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(h));
    return y;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

template <bool F16>
__global__ void __launch_bounds__(ATT_THREADS, 1)
    attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnKernelParams p) {
    // largest allowed gap between the true running max and the one the exponent still uses (log2 units):
    // the fp16 exponent argument keeps 0.1 % accuracy up to |x| = 4; the fp32 path can afford 2^8
    constexpr float TAU = F16 ? 4.0f : 8.0f;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
    uint8_t* sQ = smem;                                  // tile t at t*TILE
    uint8_t* sKV = sQ + ATT_QT * ATT_TILE;               // stage s: K at s*2*TILE, V at s*2*TILE + TILE
    uint8_t* sP = sKV + ATT_STAGES * 2 * ATT_TILE;       // tile t at t*2*TILE; two 64-wide kv chunks each
    uint8_t* sOnes = sP + ATT_QT * 2 * ATT_TILE;         // chunk c at c*2048: [16 rows x 128 B] of 1.0
    uint64_t* bars = reinterpret_cast<uint64_t*>(sOnes + ATT_ONES);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;
    uint64_t* kv_empty = kv_full + ATT_STAGES;
    uint64_t* s_full = kv_empty + ATT_STAGES;  // [QT] UMMA -> softmax : S tile ready
    uint64_t* s_free = s_full + ATT_QT;        // [QT] softmax -> UMMA : S tile copied to registers
    uint64_t* p_full = s_free + ATT_QT;        // [QT] softmax -> UMMA : P tile written (and O, L rescaled)
    uint64_t* pv_done = p_full + ATT_QT;       // [QT] UMMA -> softmax : O += P V, L += P 1 finished
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + ATT_QT);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * (ATT_BM * ATT_QT);
    const int head = blockIdx.y, b = blockIdx.z;
    const int n_tiles = p.n_kv_tiles;
    const bool has_b = q0 + ATT_BM < p.sq;  // second Q tile holds at least one live row

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(q_full, 1);
        for (int s = 0; s < ATT_STAGES; ++s) mbar_init(&kv_full[s], 1), mbar_init(&kv_empty[s], 1);
        for (int t = 0; t < ATT_QT; ++t) {
            mbar_init(&s_full[t], 1);
            mbar_init(&s_free[t], 4);
            mbar_init(&p_full[t], 4);
            mbar_init(&pv_done[t], 1);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    if (warp >= 2) {
        // constant B operand of the row-sum MMA: every element 1.0 (so the swizzle pattern is immaterial)
        const uint32_t one2 = F16 ? 0x3C003C00u : 0x3F803F80u;
        uint4* o = reinterpret_cast<uint4*>(sOnes);
        for (int i = threadIdx.x - 64; i < ATT_ONES / 16; i += ATT_THREADS - 64) o[i] = make_uint4(one2, one2, one2, one2);
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(q_full, (has_b ? 2 : 1) * ATT_TILE);
            tma_load_2d(sQ, &tmQ, q_full, head * ATT_D, b * p.sq + q0);
            if (has_b) tma_load_2d(sQ + ATT_TILE, &tmQ, q_full, head * ATT_D, b * p.sq + q0 + ATT_BM);
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait(&kv_empty[stage], phase ^ 1);
                uint8_t* sk = sKV + stage * 2 * ATT_TILE;
                mbar_expect_tx(&kv_full[stage], 2 * ATT_TILE);
                tma_load_2d(sk, &tmK, &kv_full[stage], head * ATT_D, b * p.skv + j * ATT_BN);
                tma_load_2d(sk + ATT_TILE, &tmV, &kv_full[stage], head * ATT_D, b * p.skv + j * ATT_BN);
                if (++stage == ATT_STAGES) stage = 0, phase ^= 1;
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t bf = F16 ? 0u : 1u;
            constexpr uint32_t idesc_qk = umma_idesc(128, ATT_BN, bf, bf, 0, 0);
            constexpr uint32_t idesc_pv = umma_idesc(128, ATT_D, bf, bf, 0, 1);  // V is MN-major
            constexpr uint32_t idesc_l = umma_idesc(128, 16, bf, bf, 0, 0);
            const int nq = has_b ? 2 : 1;
            const uint32_t ones_addr = smem_u32(sOnes);
            auto issue_qk = [&](int t, int stage) {
                const uint32_t q_addr = smem_u32(sQ + t * ATT_TILE);
                const uint32_t k_addr = smem_u32(sKV + stage * 2 * ATT_TILE);
                const uint32_t d = tmem_base + ATT_TMEM_S + t * 128;
#pragma unroll
                for (int ks = 0; ks < ATT_D / 16; ++ks)
                    umma_bf16(d, umma_desc_sw128(q_addr + ks * 32), umma_desc_sw128(k_addr + ks * 32), idesc_qk,
                              ks != 0);
                umma_commit(&s_full[t]);
            };
            auto issue_pv = [&](int t, int j, int stage) {
                const uint32_t v_addr = smem_u32(sKV + stage * 2 * ATT_TILE + ATT_TILE);
                const uint32_t p_addr = smem_u32(sP + t * 2 * ATT_TILE);
#pragma unroll
                for (int ks = 0; ks < ATT_BN / 16; ++ks) {
                    const uint64_t adesc = umma_desc_sw128(p_addr + (ks >> 2) * ATT_TILE + (ks & 3) * 32);
                    const uint64_t bdesc = umma_desc_sw128(v_addr + ks * 2048, /*lbo*/ 128 * 128, /*sbo*/ 1024);
                    umma_bf16(tmem_base + ATT_TMEM_O + t * ATT_D, adesc, bdesc, idesc_pv, (j | ks) != 0);
                }
#pragma unroll
                for (int ks = 0; ks < ATT_BN / 16; ++ks) {
                    const uint64_t adesc = umma_desc_sw128(p_addr + (ks >> 2) * ATT_TILE + (ks & 3) * 32);
                    const uint64_t bdesc = umma_desc_sw128(ones_addr + (ks >> 2) * 2048 + (ks & 3) * 32);
                    umma_bf16(tmem_base + ATT_TMEM_L + t * 16, adesc, bdesc, idesc_l, (j | ks) != 0);
                }
                umma_commit(&pv_done[t]);
            };
            // Event-driven issue: the two Q tiles advance independently.  Whatever is ready goes next, P V first (the
            // softmax group of that tile needs its completion before it may overwrite P), then the next Q K^T.
            mbar_wait(q_full, 0);
            tc_fence_after();
            int next_qk[ATT_QT] = {0, 0}, next_pv[ATT_QT] = {0, 0};
            int kv_ready = 0;  // kv tiles [0, kv_ready) have landed in shared memory
            int released = 0;  // kv tiles [0, released) have been handed back to the producer
            int pending = nq * n_tiles;
            while (pending > 0) {
                while (kv_ready < n_tiles &&
                       mbar_try_wait(&kv_full[kv_ready % ATT_STAGES], (kv_ready / ATT_STAGES) & 1))
                    ++kv_ready;
#pragma unroll
                for (int t = 0; t < ATT_QT; ++t) {
                    if (t >= nq) continue;
                    int j = next_pv[t];
                    if (j < next_qk[t] && mbar_try_wait(&p_full[t], j & 1)) {
                        tc_fence_after();
                        issue_pv(t, j, j % ATT_STAGES);
                        next_pv[t] = j + 1;
                        --pending;
                        int mn = next_pv[0];
                        if (nq > 1 && next_pv[1] < mn) mn = next_pv[1];
                        for (; released < mn; ++released) umma_commit(&kv_empty[released % ATT_STAGES]);
                    }
                    j = next_qk[t];
                    if (j < kv_ready && (j == 0 || mbar_try_wait(&s_free[t], (j - 1) & 1))) {
                        tc_fence_after();
                        issue_qk(t, j % ATT_STAGES);
                        next_qk[t] = j + 1;
                    }
                }
            }
        }
        __syncwarp();
    } else {
        // ---------------------------------------------------------------- softmax + output, group t = (warp-2)/4
        const int t = (warp - 2) >> 2;
        if (t == 0 || has_b) {
            const int quad = warp & 3;
            const int row = quad * 32 + lane;
            const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
            const uint32_t s_addr = tmem_base + ATT_TMEM_S + t * 128 + lane_off;
            const uint32_t o_addr = tmem_base + ATT_TMEM_O + t * ATT_D + lane_off;
            const uint32_t l_addr = tmem_base + ATT_TMEM_L + t * 16 + lane_off;
            uint8_t* prow = sP + t * 2 * ATT_TILE + row * 128;
            const float sc = p.scale_log2;
            float m_used = 0.f;
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait(&s_full[t], j & 1);
                tc_fence_after();
                uint32_t s[4][32];
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld_x32(s_addr + c * 32, s[c]);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[t]);

                const int valid = p.skv - j * ATT_BN;  // live kv columns of this tile (>= 128: all)
                if (valid < ATT_BN) {
                    // ragged last tile: dead columns get a score whose exponential is exactly 0
#pragma unroll
                    for (int c = 0; c < 4; ++c)
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c * 32 + i >= valid) s[c][i] = __float_as_uint(-INFINITY);
                }
                float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        mx0 = fmax3(mx0, __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
                        mx1 = fmax3(mx1, __uint_as_float(s[c][i + 2]), __uint_as_float(s[c][i + 3]));
                    }
                const float m_new = fmaxf(mx0, mx1) * sc;
                float alpha = 1.f;
                bool need = false;
                if (j == 0) {
                    m_used = m_new;
                } else if (m_new > m_used + TAU) {
                    alpha = ex2_ftz(m_used - m_new);
                    m_used = m_new;
                    need = true;
                }
                // exponentials first (registers only); they hide the tail of PV(j-1)
                const float neg_m = -m_used;
                uint32_t pk[16][4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            // -inf * sc stays -inf (sc > 0); the fp16 conversion saturates it and ex2 returns 0
                            const float x0 = fmaf(__uint_as_float(s[c][g * 8 + 2 * i]), sc, neg_m);
                            const float x1 = fmaf(__uint_as_float(s[c][g * 8 + 2 * i + 1]), sc, neg_m);
                            if constexpr (F16)
                                pk[c * 4 + g][i] = ex2_f16x2(x0, x1);
                            else
                                pk[c * 4 + g][i] = pack_bf16x2(ex2_ftz(x0), ex2_ftz(x1));
                        }
                    }
                }
                // PV(j-1) must be complete before P_t is overwritten or O_t / L_t rescaled
                if (j > 0) {
                    mbar_wait(&pv_done[t], (j - 1) & 1);
                    tc_fence_after();
                }
                if (__any_sync(0xffffffffu, need)) {
#pragma unroll
                    for (int c = 0; c < ATT_D / 16 + 1; ++c) {  // four O chunks, then the L chunk
                        const uint32_t a = c < ATT_D / 16 ? o_addr + c * 16 : l_addr;
                        uint32_t o[16];
                        tmem_ld_x16(a, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st_x16(a, o);
                    }
                    tmem_st_wait();
                }
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const int col0 = q * 8;
                    const int chunk = col0 >> 6;      // which 64-wide kv chunk
                    const int cc = (col0 & 63) >> 3;  // 16-byte chunk inside the 128-byte row
                    *reinterpret_cast<uint4*>(prow + chunk * ATT_TILE + ((cc ^ (row & 7)) << 4)) =
                        make_uint4(pk[q][0], pk[q][1], pk[q][2], pk[q][3]);
                }
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[t]);
            }
            // ---- epilogue: O / L -> 16-bit
            mbar_wait(&pv_done[t], (n_tiles - 1) & 1);
            tc_fence_after();
            float inv;
            {
                uint32_t lv[16];
                tmem_ld_x16(l_addr, lv);
                tmem_ld_wait();
                inv = 1.0f / __uint_as_float(lv[0]);
            }
            const int qrow = q0 + t * ATT_BM + row;
            uint16_t* orow = p.out + ((long long)b * p.sq + qrow) * p.ld_out + head * ATT_D;
#pragma unroll
            for (int c = 0; c < ATT_D / 16; ++c) {
                uint32_t o[16];
                tmem_ld_x16(o_addr + c * 16, o);
                tmem_ld_wait();
                if (qrow < p.sq) {
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        pk[i] = pack_h2(__uint_as_float(o[2 * i]) * inv, __uint_as_float(o[2 * i + 1]) * inv, F16);
                    *reinterpret_cast<uint4*>(orow + c * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4*>(orow + c * 16 + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

template <bool F16>
static int32_t launch_attention(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                const AttnKernelParams& kp, dim3 grid, cudaStream_t stream) {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t e =
            cudaFuncSetAttribute(attention_kernel<F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM);
        if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "attention smem attribute: %s", cudaGetErrorString(e));
        attr_set[dev] = true;
    }
    attention_kernel<F16><<<grid, ATT_THREADS, ATT_SMEM, stream>>>(tmQ, tmK, tmV, kp);
    count_launch();
    B200SD_CHECK_LAUNCH("attention_kernel");
    return B200SD_OK;
}

}  // namespace b200sd

using namespace b200sd;

extern "C" int32_t b200sd_attention(const b200sd_attention_params* p, b200sd_stream_t stream) {
    B200SD_CHECK_ARG(p && p->q && p->k && p->v && p->out, "attention: null pointer");
    if (p->head_dim != 64) return set_error(B200SD_ERR_UNSUPPORTED, "attention: head_dim %d (only 64)", p->head_dim);
    B200SD_CHECK_ARG(p->batch > 0 && p->heads > 0 && p->sq > 0 && p->skv > 0, "attention: bad shape");
    B200SD_CHECK_ARG(p->dtype == B200SD_BF16 || p->dtype == B200SD_F16, "attention: dtype must be bf16 or fp16");
    B200SD_CHECK_ARG(p->scale > 0.f, "attention: scale must be positive");
    B200SD_CHECK_ARG(p->ld_q % 8 == 0 && p->ld_k % 8 == 0 && p->ld_v % 8 == 0 && p->ld_out % 8 == 0,
                     "attention: leading dimensions must be multiples of 8");
    B200SD_CHECK_ARG(((reinterpret_cast<uintptr_t>(p->q) | reinterpret_cast<uintptr_t>(p->k) |
                       reinterpret_cast<uintptr_t>(p->v) | reinterpret_cast<uintptr_t>(p->out)) & 15) == 0,
                     "attention: pointers must be 16-byte aligned");
    B200SD_CHECK_ARG(p->batch <= 65535 && p->heads <= 65535, "attention: batch/heads too large for the grid");
    CUtensorMap tmQ, tmK, tmV;
    const uint32_t box[2] = {ATT_D, 128};
    {
        uint64_t dims[2] = {(uint64_t)p->heads * ATT_D, (uint64_t)p->batch * p->sq};
        uint64_t str[1] = {(uint64_t)p->ld_q * 2};
        int32_t r = encode_tmap_bf16(&tmQ, p->q, 2, dims, str, box);
        if (r) return r;
    }
    {
        uint64_t dims[2] = {(uint64_t)p->heads * ATT_D, (uint64_t)p->batch * p->skv};
        uint64_t strk[1] = {(uint64_t)p->ld_k * 2};
        uint64_t strv[1] = {(uint64_t)p->ld_v * 2};
        int32_t r = encode_tmap_bf16(&tmK, p->k, 2, dims, strk, box);
        if (r) return r;
        r = encode_tmap_bf16(&tmV, p->v, 2, dims, strv, box);
        if (r) return r;
    }
    AttnKernelParams kp;
    kp.sq = p->sq, kp.skv = p->skv;
    kp.n_kv_tiles = (p->skv + ATT_BN - 1) / ATT_BN;
    kp.scale_log2 = p->scale * 1.4426950408889634f;
    kp.out = static_cast<uint16_t*>(p->out);
    kp.ld_out = p->ld_out;
    const int rows_per_cta = ATT_BM * ATT_QT;
    dim3 grid((p->sq + rows_per_cta - 1) / rows_per_cta, p->heads, p->batch);
    if (p->dtype == B200SD_F16) return launch_attention<true>(tmQ, tmK, tmV, kp, grid, as_stream(stream));
    return launch_attention<false>(tmQ, tmK, tmV, kp, grid, as_stream(stream));
}
