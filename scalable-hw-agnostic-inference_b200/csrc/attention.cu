// K3: fused flash attention for head_dim 64 on sm_100a (self- and cross-attention of BasicTransformerBlock).
//
// One CTA owns 256 query rows = two 128-row tiles (A, B) that share every K/V tile:
//   warp 0        TMA producer : both Q tiles once, then K/V tiles (128 kv rows x 64) through a 4-stage ring
//   warp 1, 18    UMMA issuers : one warp per Q tile.  S_t = Q_t K^T (128x128x64, both operands from shared memory) into
//                                TMEM; O_t += P_t V (128x64x128) with P_t read FROM TMEM (tcgen05.mma A-in-TMEM form) and V
//                                as an MN-major shared-memory operand -- no transpose of V and no shared-memory round trip
//                                for P.  A tile's events come in a fixed order, so its warp issues in order behind
//                                blocking barrier waits (Q K^T of step j+1, then P V of step j).
//   warps 2..9    softmax A    } either two threads per query row, 64 score columns each (tcgen05.ld 32x32b; one row-max
//   warps 10..17  softmax B    } exchange per tile), or -- default -- a row spread over the four threads of a quad
//                                (tcgen05.ld 16x256b; row max = two shuffles, nothing crosses warps);
//                                the scores are held in registers so S_t is released to the next QK^T at once; exp2 with
//                                the scale folded in; lazy rescale of O (only when the running max grows by more than
//                                2^8); P packed to 16 bits and written back to TMEM with one tcgen05.st per thread.
// Why P lives in TMEM: measured with P staged in shared memory, the kernel time did not move when the softmax work
// was halved or the softmax warps doubled -- the shared-memory port (Q, K, V, P reads by the tensor core + P writes)
// and the UMMA issue side were the bound, not the exponentials.
// The S x S score matrix never leaves the SM.  FLOPs per launch: 4 * batch * heads * sq * skv * 64.
#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace b200sd {
using namespace ptx;

constexpr int ATT_BM = 128;          // rows per Q tile
constexpr int ATT_QT = 2;            // Q tiles per CTA
constexpr int ATT_BN = 128;          // kv rows per tile
constexpr int ATT_D = 64;
constexpr int ATT_STAGES = 4;
constexpr int ATT_TILE = 128 * 128;  // bytes of a [128 x 64] 16-bit tile
constexpr int ATT_THREADS = 32 * (3 + 8 * ATT_QT);  // TMA, UMMA (tile A), 8 softmax warps per Q tile, UMMA (tile B)
constexpr int ATT_MMA_B_WARP = 2 + 8 * ATT_QT;
constexpr int ATT_XCH = 3 * ATT_QT * 256 * 4;       // row-max exchange [parity][tile][half][128] fp32 + row-sum exchange [tile][half][128]
constexpr int ATT_SMEM = ATT_QT * ATT_TILE /*Q*/ + ATT_STAGES * 2 * ATT_TILE /*K,V*/ + ATT_XCH + 1024 + 256;
// TMEM columns: S_A 0..127, S_B 128..255, O_A 256..319, O_B 320..383, P_A 384..447, P_B 448..511 (16-bit pairs)
constexpr uint32_t ATT_TMEM_S = 0, ATT_TMEM_O = 256, ATT_TMEM_P = 384;

struct AttnKernelParams {
    int sq, skv, n_kv_tiles;
    int heads, n_qpairs, items;  // work items = (batch, head, pair of 128-row Q tiles), walked by persistent CTAs
    float scale_log2;  // scale * log2(e)
    uint16_t* out;
    long long ld_out;
    int causal;   // 1: key j is visible to query i only if j <= i
    int stagger;  // 1: Q tile B starts half a softmax period behind tile A (see the UMMA warp)
    int pair_bar; // (half-row layout) 1: the two warps that share a row synchronise alone (64-thread barrier) instead of per tile
};

__device__ __forceinline__ float ex2_ftz(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// 2^x on the FMA / ALU pipes (no MUFU): round-to-nearest split x = n + f, degree-3 minimax polynomial for 2^f on
// [-0.5, 0.5] (max relative error 7.5e-5, well under half an fp16 ulp), then n added into the exponent field.
// Valid for x <= 64; anything below -30 (including -inf) returns a value that rounds to 0 in 16 bits.
__device__ __forceinline__ float exp2_poly(float x) {
    x = fmaxf(x, -30.f);
    const float t = x + 12582912.f;  // 1.5 * 2^23: the low mantissa bits of t now hold round(x)
    const float f = x - (t - 12582912.f);
    const float p = fmaf(fmaf(fmaf(0.055171668f, f, 0.24261112f), f, 0.69326099f), f, 0.99992807f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// exp2_poly on a pair, with the additions and the polynomial on FFMA2 / FADD2: 5 issue slots per value, no MUFU
__device__ __forceinline__ void exp2_poly2(uint64_t x2, float& e0, float& e1) {
    float x0, x1;
    unpack2(x2, x0, x1);
    x2 = pack2(fmaxf(x0, -30.f), fmaxf(x1, -30.f));
    const uint64_t magic = pack2(12582912.f, 12582912.f), nmagic = pack2(-12582912.f, -12582912.f);
    const uint64_t t2 = add2(x2, magic);                       // low mantissa bits = round(x)
    const uint64_t f2 = fma2(add2(t2, nmagic), pack2(-1.f, -1.f), x2);  // x - round(x) in [-0.5, 0.5]
    uint64_t p2 = fma2(pack2(0.055171668f, 0.055171668f), f2, pack2(0.24261112f, 0.24261112f));
    p2 = fma2(p2, f2, pack2(0.69326099f, 0.69326099f));
    p2 = fma2(p2, f2, pack2(0.99992807f, 0.99992807f));
    float p0, p1, t0, t1;
    unpack2(p2, p0, p1);
    unpack2(t2, t0, t1);
    e0 = __int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23));
    e1 = __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23));
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_st_x32(uint32_t taddr, const uint32_t (&r)[32]) {
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

// EMU_MASK: bit i set = pair i of every eight pairs of exponentials is evaluated by exp2_poly2 instead of the MUFU.
// QUADROWS: the softmax warps read S with the 16-lane TMEM shapes (a row lives in the four threads of a quad, a warp owns
// 16 whole rows) instead of one half-row per thread: no row-max exchange between warps, no named barrier per kv step.
template <bool F16, int EMU_MASK, bool PERSIST, bool QUADROWS>
__global__ void __launch_bounds__(ATT_THREADS, 1)
    attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnKernelParams p) {
    constexpr float TAU = 8.0f;  // largest lag (log2 units) of the max the exponent uses behind the true running max
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
    uint8_t* sQ = smem;                                  // tile t at t*TILE
    uint8_t* sKV = sQ + ATT_QT * ATT_TILE;               // stage s: K at s*2*TILE, V at s*2*TILE + TILE
    float* xchg = reinterpret_cast<float*>(sKV + ATT_STAGES * 2 * ATT_TILE);  // [parity][tile][half][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(xchg) + ATT_XCH);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;
    uint64_t* kv_empty = kv_full + ATT_STAGES;
    uint64_t* s_full = kv_empty + ATT_STAGES;  // [QT] UMMA -> softmax : S tile ready
    uint64_t* s_free = s_full + ATT_QT;        // [QT] softmax -> UMMA : S tile copied to registers
    uint64_t* p_full = s_free + ATT_QT;        // [QT] softmax -> UMMA : P tile in TMEM (and O rescaled)
    uint64_t* pv_done = p_full + ATT_QT;       // [QT] UMMA -> softmax : O += P V finished
    uint64_t* b_go = pv_done + ATT_QT;         // softmax(A) -> UMMA : tile A entered its first exponential phase
    uint64_t* q_empty = b_go + 1;              // UMMA -> TMA : every Q K^T of the work item has read the Q tiles
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = p.n_kv_tiles;
    // Persistent CTAs: work item `it` = (batch b, head, pair of Q tiles at row q0).  Short problems (cross-attention with
    // 77 keys, the 24x24 level) are dominated by per-CTA set-up -- barrier init, TMEM allocation, the first TMA round
    // trip -- and one CTA per SM cannot overlap two of them, so a CTA keeps its barriers and TMEM and walks several items;
    // the Q tiles of the next item are requested as soon as the last Q K^T of the current one has been issued.
    auto item_q0 = [&](int it) { return (it % p.n_qpairs) * (ATT_BM * ATT_QT); };
    auto item_head = [&](int it) { return (it / p.n_qpairs) % p.heads; };
    auto item_b = [&](int it) { return it / (p.n_qpairs * p.heads); };

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(q_full, 1);
        for (int s = 0; s < ATT_STAGES; ++s) mbar_init(&kv_full[s], 1), mbar_init(&kv_empty[s], ATT_QT);  // one report per issuing warp
        for (int t = 0; t < ATT_QT; ++t) {
            mbar_init(&s_full[t], 1);
            mbar_init(&s_free[t], 8);
            mbar_init(&p_full[t], 8);
            mbar_init(&pv_done[t], 1);
        }
        mbar_init(b_go, 8);
        mbar_init(q_empty, ATT_QT);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_sync();  // set-up above overlaps the previous kernel's tail; Q/K/V are read and O written from here on

    // Warps 0 and 1 keep all 32 lanes in warp-uniform control flow and predicate only the asynchronous instructions
    // on one elected lane; barrier probes are made uniform with a vote.  Issued from inside an `if (lane == 0)`
    // region, every UMMA / TMA operand is thread-varying to the compiler and each instruction gets wrapped in an
    // R2UR + ELECT + BRA.U.ANY waterfall loop: measured, ~700 scalar instructions per kv step at ~5 cycles each on
    // the one issuing thread -- the whole kernel ran at the speed of that thread.
    constexpr uint32_t FULL = 0xffffffffu;
    if (warp == 0) {
        int stage = 0;
        uint32_t phase = 0;
        int n_it = 0;
        for (int it = blockIdx.x; it < p.items; it += (PERSIST ? (int)gridDim.x : p.items), ++n_it) {
            const int q0 = item_q0(it), head = item_head(it), b = item_b(it);
            const bool has_b = q0 + ATT_BM < p.sq;  // second Q tile holds at least one live row
            if (PERSIST && n_it > 0) mbar_wait(q_empty, (n_it - 1) & 1);
            if (elect_one()) {
                mbar_expect_tx(q_full, (has_b ? 2 : 1) * ATT_TILE);
                tma_load_2d(sQ, &tmQ, q_full, head * ATT_D, b * p.sq + q0);
                if (has_b) tma_load_2d(sQ + ATT_TILE, &tmQ, q_full, head * ATT_D, b * p.sq + q0 + ATT_BM);
            }
            __syncwarp();
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait(&kv_empty[stage], phase ^ 1);
                uint8_t* sk = sKV + stage * 2 * ATT_TILE;
                if (elect_one()) {
                    mbar_expect_tx(&kv_full[stage], 2 * ATT_TILE);
                    tma_load_2d(sk, &tmK, &kv_full[stage], head * ATT_D, b * p.skv + j * ATT_BN);
                    tma_load_2d(sk + ATT_TILE, &tmV, &kv_full[stage], head * ATT_D, b * p.skv + j * ATT_BN);
                }
                __syncwarp();
                if (++stage == ATT_STAGES) stage = 0, phase ^= 1;
            }
        }
    } else if (warp == 1 || warp == ATT_MMA_B_WARP) {
        // ---------------------------------------------------------------- UMMA issue: one warp per Q tile
        // Each Q tile has its own issuing warp.  For ONE tile the events come in a fixed order -- S(j) read by the
        // softmax group (s_free) always precedes P(j) written (p_full) -- so the warp waits for them in that order with
        // blocking waits and needs no polling.  Measured with one warp polling for both tiles (ncu source view, S = 9216):
        // that warp's loop was ~600 dependent scalar instructions per kv step, ~3000 cycles, and it found all four of
        // {P_A V, Q_A K^T, P_B V, Q_B K^T} ready on every pass: the issuing warp, not the exponentials or the tensor pipe
        // (34 % active), set the kernel's speed, and the softmax warps spent 46 % of their time waiting for it.
        constexpr uint32_t bf = F16 ? 0u : 1u;
        constexpr uint32_t idesc_qk = umma_idesc(128, ATT_BN, bf, bf, 0, 0);
        constexpr uint32_t idesc_pv = umma_idesc(128, ATT_D, bf, bf, 0, 1);  // A = P (TMEM, K-major), B = V MN-major
        const int t = warp == 1 ? 0 : 1;
        const uint32_t tmem_u = __shfl_sync(FULL, tmem_base, 0);
        const uint32_t q_u = __shfl_sync(FULL, smem_u32(sQ), 0);
        const uint32_t kv_u = __shfl_sync(FULL, smem_u32(sKV), 0);
        // descriptors of stage 0; stage s adds s * (2 * TILE >> 4) to the start-address field (no carry: the field holds
        // addr >> 4 of a 227 KB window in 14 bits)
        const uint64_t qdesc = umma_desc_sw128(q_u + t * ATT_TILE);
        const uint64_t kdesc0 = umma_desc_sw128(kv_u);
        const uint64_t vdesc0 = umma_desc_sw128(kv_u + ATT_TILE, /*lbo*/ 128 * 128, /*sbo*/ 1024);
        constexpr uint32_t STAGE_STEP = (2 * ATT_TILE) >> 4;
        const uint32_t s_tmem = tmem_u + ATT_TMEM_S + t * 128;
        const uint32_t p_tmem = tmem_u + ATT_TMEM_P + t * 64;
        const uint32_t o_tmem = tmem_u + ATT_TMEM_O + t * ATT_D;
        auto issue_qk = [&](int stage) {
            const uint64_t kdesc = kdesc0 + (uint64_t)(stage * STAGE_STEP);
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < ATT_D / 16; ++ks)  // +32 bytes along K = +2 in the (addr >> 4) field
                    umma_bf16(s_tmem, qdesc + 2 * ks, kdesc + 2 * ks, idesc_qk, ks != 0);
                umma_commit(&s_full[t]);
            }
            __syncwarp();
        };
        auto issue_pv = [&](int j, int stage) {
            const uint64_t vdesc = vdesc0 + (uint64_t)(stage * STAGE_STEP);
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < ATT_BN / 16; ++ks)
                    // 16 K-elements of P = 8 TMEM columns (two 16-bit values per column); 16 kv rows of V = 2048 bytes
                    umma_ts(o_tmem, p_tmem + ks * 8, vdesc + (uint64_t)(ks * (2048 >> 4)), idesc_pv, (j | ks) != 0);
                umma_commit(&pv_done[t]);
                umma_commit(&kv_empty[stage]);  // this tile is done with the stage (the other tile's warp reports too)
            }
            __syncwarp();
        };
        // Barrier parities run over the whole life of the CTA: gk = kv tiles of all earlier items (the K/V ring),
        // gt = kv steps this Q tile has done before this item (tile B sits out items whose second tile is empty).
        int gk = 0, gt = 0, n_it = 0;
        for (int it = blockIdx.x; it < p.items; it += (PERSIST ? (int)gridDim.x : p.items), ++n_it) {
            const bool active = t == 0 || item_q0(it) + ATT_BM < p.sq;
            const int NI = PERSIST ? n_it : 0, GK = PERSIST ? gk : 0, GT = PERSIST ? gt : 0;
            auto kv_stage = [&](int j) { return (GK + j) % ATT_STAGES; };
            auto kv_parity = [&](int j) { return (uint32_t)(((GK + j) / ATT_STAGES) & 1); };
            if (!active) {
                // this tile holds no live row: the ring and the Q buffer still expect one report per issuing warp.  It also
                // waits on the once-per-item barriers like an active tile would, so that it can never run a whole phase
                // ahead of them (a parity wait cannot tell phase n from phase n - 2)
                mbar_wait(q_full, NI & 1);
                if (p.stagger) mbar_wait(b_go, NI & 1);
                for (int j = 0; j < n_tiles; ++j) {
                    mbar_wait(&kv_full[kv_stage(j)], kv_parity(j));
                    if (elect_one()) mbar_arrive(&kv_empty[kv_stage(j)]);
                    __syncwarp();
                }
                if (PERSIST) {
                    if (elect_one()) mbar_arrive(q_empty);
                    __syncwarp();
                }
            } else {
                mbar_wait(q_full, NI & 1);
                // The softmax groups of the two tiles alternate between a MUFU-bound phase (exponentials) and a phase that
                // does not touch the MUFU (TMEM loads, row max, P store).  Started together they stay in step and fight
                // over the MUFU, then leave it idle together; tile B's first Q K^T therefore waits until tile A has
                // entered its first exponential phase, and the offset persists (the loops are self-paced).
                if (t == 1 && p.stagger) mbar_wait(b_go, NI & 1);
                mbar_wait(&kv_full[kv_stage(0)], kv_parity(0));
                // the S tile must have been read by the softmax group (the last step of this tile's previous item)
                if (GT > 0) mbar_wait(&s_free[t], (GT - 1) & 1);
                tc_fence_after();
                issue_qk(kv_stage(0));
                if (PERSIST && n_tiles == 1) {
                    if (elect_one()) umma_commit(q_empty);
                    __syncwarp();
                }
                for (int j = 0; j < n_tiles; ++j) {
                    if (j + 1 < n_tiles) {
                        mbar_wait(&kv_full[kv_stage(j + 1)], kv_parity(j + 1));
                        mbar_wait(&s_free[t], (GT + j) & 1);
                        tc_fence_after();
                        issue_qk(kv_stage(j + 1));
                        // every Q K^T of this item is issued: once they complete, the next item's Q tiles may land
                        if (PERSIST && j + 2 == n_tiles) {
                            if (elect_one()) umma_commit(q_empty);
                            __syncwarp();
                        }
                    }
                    mbar_wait(&p_full[t], (GT + j) & 1);
                    tc_fence_after();
                    issue_pv(j, kv_stage(j));
                }
                gt += n_tiles;
            }
            gk += n_tiles;
        }
    } else if (QUADROWS) {
        // ---------------------------------------------------------------- softmax + output, quad-row layout
        // 8 warps per Q tile; warp (quad, half) owns rows quad*32 + half*16 .. +15 of the tile and ALL 128 score columns
        // of them: thread t holds rows t/4 and t/4 + 8, columns 8i + 2(t%4), +1 for i = 0..15 (tcgen05.ld 16x256b).
        // The row max is two shuffles inside the quad; nothing is exchanged between warps, so the warps of a tile drift
        // apart freely and spread their MUFU phases instead of hitting the pipe in lock-step after a barrier.
        const int t = (warp - 2) >> 3;
        int gt = 0, n_it = 0;
        for (int it = blockIdx.x; it < p.items; it += (PERSIST ? (int)gridDim.x : p.items), ++n_it) {
        const int q0 = item_q0(it), head = item_head(it), b = item_b(it);
        const bool has_b = q0 + ATT_BM < p.sq;
        const int GTS = PERSIST ? gt : 0;
        if (t == 0 || has_b) {
            const int half = ((warp - 2) >> 2) & 1;
            const int quad = warp & 3;
            const int wrow0 = quad * 32 + half * 16;   // first tile row of this warp
            const int r_in = lane >> 2, cq = lane & 3;  // thread: rows wrow0 + r_in, + 8; column pair index cq
            const uint32_t lane_off = (uint32_t)wrow0 << 16;
            const uint32_t s_addr = tmem_base + ATT_TMEM_S + t * 128 + lane_off;
            const uint32_t o_addr = tmem_base + ATT_TMEM_O + t * ATT_D + lane_off;
            const uint32_t p_addr = tmem_base + ATT_TMEM_P + t * 64 + lane_off;
            const float sc = p.scale_log2;
            float m_used[2] = {0.f, 0.f}, l[2] = {0.f, 0.f};
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait(&s_full[t], (GTS + j) & 1);
                tc_fence_after();
                uint32_t s[64];
                tmem_ld_16x256b_x16(s_addr, s);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[t]);

                // masking stays behind a warp-uniform branch (see the half-row layout below)
                const int valid = p.skv - j * ATT_BN;  // live kv columns of this tile
                const int kbase = j * ATT_BN;
                const bool diag = p.causal && (kbase + ATT_BN - 1 > q0 + t * ATT_BM + wrow0);
                if (valid < ATT_BN || diag) {
                    __syncwarp();
                    const int qa = p.causal ? q0 + t * ATT_BM + wrow0 + r_in : 0x7fffffff;
                    const int qb = p.causal ? qa + 8 : 0x7fffffff;
#pragma unroll
                    for (int i = 0; i < 16; ++i)
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int col = 8 * i + 2 * cq + (e & 1);
                            if (col >= valid || kbase + col > ((e & 2) ? qb : qa)) s[4 * i + e] = __float_as_uint(-INFINITY);
                        }
                }
                float mxa = -INFINITY, mxb = -INFINITY;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    mxa = fmax3(mxa, __uint_as_float(s[4 * i]), __uint_as_float(s[4 * i + 1]));
                    mxb = fmax3(mxb, __uint_as_float(s[4 * i + 2]), __uint_as_float(s[4 * i + 3]));
                }
                mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, 1));
                mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, 1));
                mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, 2));
                mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, 2));
                if (t == 0 && j == 0 && lane == 0) mbar_arrive(b_go);
                const float m_new[2] = {mxa * sc, mxb * sc};
                float alpha[2] = {1.f, 1.f};
                bool need = false;
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    if (j == 0) {
                        m_used[r] = m_new[r];
                    } else if (m_new[r] > m_used[r] + TAU) {
                        alpha[r] = ex2_ftz(m_used[r] - m_new[r]);
                        m_used[r] = m_new[r];
                        need = true;
                    }
                }
                uint32_t pk[32];
                const uint64_t sc2 = pack2(sc, sc);
                const uint64_t negm2[2] = {pack2(-m_used[0], -m_used[0]), pack2(-m_used[1], -m_used[1])};
                uint64_t ls2[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};
#pragma unroll
                for (int i = 0; i < 16; ++i) {
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        const uint64_t x2 = fma2(pack2(__uint_as_float(s[4 * i + 2 * r]), __uint_as_float(s[4 * i + 2 * r + 1])),
                                                 sc2, negm2[r]);
                        float e0, e1;
                        if ((EMU_MASK >> ((2 * i + r) & 7)) & 1) {
                            exp2_poly2(x2, e0, e1);
                        } else {
                            float x0, x1;
                            unpack2(x2, x0, x1);
                            e0 = ex2_ftz(x0), e1 = ex2_ftz(x1);
                        }
                        ls2[r] = add2(ls2[r], pack2(e0, e1));
                        pk[2 * i + r] = pack_h2(e0, e1, F16);
                    }
                }
                if (j > 0) {
                    mbar_wait(&pv_done[t], (GTS + j - 1) & 1);
                    tc_fence_after();
                }
                if (__any_sync(0xffffffffu, need)) {
                    uint32_t o[32];  // rows r_in / r_in + 8, columns 8i + 2cq, +1 of the 64 head-dim columns
                    tmem_ld_16x256b_x8(o_addr, o);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        o[4 * i] = __float_as_uint(__uint_as_float(o[4 * i]) * alpha[0]);
                        o[4 * i + 1] = __float_as_uint(__uint_as_float(o[4 * i + 1]) * alpha[0]);
                        o[4 * i + 2] = __float_as_uint(__uint_as_float(o[4 * i + 2]) * alpha[1]);
                        o[4 * i + 3] = __float_as_uint(__uint_as_float(o[4 * i + 3]) * alpha[1]);
                    }
                    tmem_st_16x256b_x8(o_addr, o);
                }
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    float a0, a1;
                    unpack2(ls2[r], a0, a1);
                    l[r] = l[r] * alpha[r] + (a0 + a1);
                }
                tmem_st_16x128b_x16(p_addr, pk);  // 128 probabilities per row = 64 TMEM columns, 16 of them from this thread
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[t]);
            }
            // ---- epilogue: O / l -> 16-bit
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                l[r] += __shfl_xor_sync(0xffffffffu, l[r], 1);
                l[r] += __shfl_xor_sync(0xffffffffu, l[r], 2);
            }
            const float inv[2] = {1.0f / l[0], 1.0f / l[1]};
            mbar_wait(&pv_done[t], (GTS + n_tiles - 1) & 1);
            tc_fence_after();
            uint32_t o[32];
            tmem_ld_16x256b_x8(o_addr, o);
            tmem_ld_wait();
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const int qrow = q0 + t * ATT_BM + wrow0 + r_in + 8 * r;
                if (qrow < p.sq) {
                    uint32_t* orow = reinterpret_cast<uint32_t*>(p.out + ((long long)b * p.sq + qrow) * p.ld_out + head * ATT_D);
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        orow[4 * i + cq] = pack_h2(__uint_as_float(o[4 * i + 2 * r]) * inv[r],
                                                   __uint_as_float(o[4 * i + 2 * r + 1]) * inv[r], F16);
                }
            }
            gt += n_tiles;
        }
        }
    } else {
        // ---------------------------------------------------------------- softmax + output
        // 8 warps per Q tile: tile t = (warp-2)/8; inside a tile warps 0..3 own score columns 0..63 of their 32 rows,
        // warps 4..7 own columns 64..127 of the same rows (TMEM lane quadrant = warp % 4 either way).
        const int t = (warp - 2) >> 3;
        int gt = 0, n_it = 0;  // kv steps this Q tile has done in earlier items; items this CTA has started
        for (int it = blockIdx.x; it < p.items; it += (PERSIST ? (int)gridDim.x : p.items), ++n_it) {
        const int q0 = item_q0(it), head = item_head(it), b = item_b(it);
        const bool has_b = q0 + ATT_BM < p.sq;
        const int GTS = PERSIST ? gt : 0;  // parity base of this tile's barriers
        if (t == 0 || has_b) {
            const int half = ((warp - 2) >> 2) & 1;
            const int quad = warp & 3;
            const int row = quad * 32 + lane;
            const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
            const uint32_t s_addr = tmem_base + ATT_TMEM_S + t * 128 + half * 64 + lane_off;
            const uint32_t o_addr = tmem_base + ATT_TMEM_O + t * ATT_D + half * 32 + lane_off;
            const uint32_t p_addr = tmem_base + ATT_TMEM_P + t * 64 + half * 32 + lane_off;
            const float sc = p.scale_log2;
            float m_used = 0.f, l = 0.f;
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait(&s_full[t], (GTS + j) & 1);
                tc_fence_after();
                uint32_t s[2][32];
#pragma unroll
                for (int c = 0; c < 2; ++c) tmem_ld_x32(s_addr + c * 32, s[c]);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[t]);

                // Masking (ragged last kv tile, causal diagonal) touches every score, so it must stay OFF the common path:
                // the guard is warp-uniform and the block holds a convergent __syncwarp, which keeps the compiler from
                // flattening it into 128 unconditional compare+select pairs (measured: they were more instructions than
                // the softmax itself).
                const int valid = p.skv - j * ATT_BN - half * 64;  // live kv columns among this thread's 64
                const int kbase = j * ATT_BN + half * 64;
                const bool diag = p.causal && (kbase + 63 > q0 + t * ATT_BM + quad * 32);  // some lane of this warp is cut
                if (valid < 64 || diag) {
                    __syncwarp();
                    const int qrow = p.causal ? q0 + t * ATT_BM + row : 0x7fffffff;
#pragma unroll
                    for (int c = 0; c < 2; ++c)
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c * 32 + i >= valid || kbase + c * 32 + i > qrow) s[c][i] = __float_as_uint(-INFINITY);
                }
                float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
                for (int c = 0; c < 2; ++c)
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        mx0 = fmax3(mx0, __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
                        mx1 = fmax3(mx1, __uint_as_float(s[c][i + 2]), __uint_as_float(s[c][i + 3]));
                    }
                // the other half of this row lives in the warp 4 further on (or back): exchange through shared memory,
                // double-buffered by tile parity so one named barrier per tile suffices
                float* xch = xchg + (((GTS + j) & 1) * ATT_QT + t) * 256;
                xch[half * 128 + row] = fmaxf(mx0, mx1);
                if (p.pair_bar) named_bar_sync(4 + t * 4 + quad, 64); else named_bar_sync(2 + t, 256);
                if (t == 0 && j == 0 && lane == 0) mbar_arrive(b_go);
                const float m_new = fmaxf(xch[row], xch[128 + row]) * sc;
                float alpha = 1.f;
                bool need = false;
                if (j == 0) {
                    m_used = m_new;
                } else if (m_new > m_used + TAU) {
                    alpha = ex2_ftz(m_used - m_new);
                    m_used = m_new;
                    need = true;
                }
                // exponentials (registers only); they hide the tail of PV(j-1)
                const float neg_m = -m_used;
                uint32_t pk[32];
                const uint64_t sc2 = pack2(sc, sc), negm2 = pack2(neg_m, neg_m);
                uint64_t ls2 = pack2(0.f, 0.f);
#pragma unroll
                for (int c = 0; c < 2; ++c) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        // -inf * sc stays -inf (sc > 0) and both exponentials return 0 for it
                        const uint64_t x2 =
                            fma2(pack2(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1])), sc2, negm2);
                        float e0, e1;
                        if ((EMU_MASK >> (i & 7)) & 1) {
                            exp2_poly2(x2, e0, e1);
                        } else {
                            float x0, x1;
                            unpack2(x2, x0, x1);
                            e0 = ex2_ftz(x0), e1 = ex2_ftz(x1);
                        }
                        ls2 = add2(ls2, pack2(e0, e1));
                        pk[c * 16 + i] = pack_h2(e0, e1, F16);
                    }
                }
                float ls0, ls1;
                unpack2(ls2, ls0, ls1);
                // PV(j-1) must be complete before P_t is overwritten or O_t rescaled
                if (j > 0) {
                    mbar_wait(&pv_done[t], (GTS + j - 1) & 1);
                    tc_fence_after();
                }
                if (__any_sync(0xffffffffu, need)) {
#pragma unroll
                    for (int c = 0; c < 2; ++c) {  // this thread's 32 O columns
                        uint32_t o[16];
                        tmem_ld_x16(o_addr + c * 16, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st_x16(o_addr + c * 16, o);
                    }
                }
                l = l * alpha + (ls0 + ls1);
                tmem_st_x32(p_addr, pk);  // 64 probabilities = 32 TMEM columns of this row
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[t]);
            }
            // ---- epilogue: O / l -> 16-bit (this thread's 32 of the 64 head-dim columns)
            float* xch = xchg + (2 * ATT_QT + t) * 256;  // its own buffer: the next item's first row max may already be
                                                         // on its way into the parity buffers
            xch[half * 128 + row] = l;
            if (p.pair_bar) named_bar_sync(4 + t * 4 + quad, 64); else named_bar_sync(2 + t, 256);
            const float inv = 1.0f / (xch[row] + xch[128 + row]);
            mbar_wait(&pv_done[t], (GTS + n_tiles - 1) & 1);
            tc_fence_after();
            const int qrow = q0 + t * ATT_BM + row;
            uint16_t* orow = p.out + ((long long)b * p.sq + qrow) * p.ld_out + head * ATT_D + half * 32;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
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
            gt += n_tiles;
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

template <bool F16, int EMU_MASK, bool PERSIST, bool QUADROWS>
static int32_t launch_attention(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                                const AttnKernelParams& kp, dim3 grid, cudaStream_t stream) {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(attention_kernel<F16, EMU_MASK, PERSIST, QUADROWS>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM);
        if (e != cudaSuccess) return set_error(B200SD_ERR_CUDA, "attention smem attribute: %s", cudaGetErrorString(e));
        attr_set[dev] = true;
    }
    B200SD_LAUNCH("attention_kernel", (attention_kernel<F16, EMU_MASK, PERSIST, QUADROWS>), grid, dim3(ATT_THREADS), ATT_SMEM, stream, tmQ,
                  tmK, tmV, kp);
    count_launch();
    B200SD_CHECK_LAUNCH("attention_kernel");
    return B200SD_OK;
}

static int stagger_from_env() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("B200SD_ATTN_STAGGER");
        cached = (e && e[0] == '0') ? 0 : 1;
    }
    return cached;
}

// softmax register layout: 0 = half rows per thread (default), 1 = quad rows (16-lane TMEM shapes, no exchange between
// warps).  Measured inside the batch-16 pipeline (power-capped clocks): half rows with the pair barrier 740 ms of attention
// per pass, quad rows 760 ms; standalone at S = 9216 both run at ~905 TFLOP/s (profiles/r02_attn_sweep.txt)
static int layout_from_env() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("B200SD_ATTN_LAYOUT");
        cached = (e && e[0] == '1') ? 1 : 0;
    }
    return cached;
}

static int pair_bar_from_env() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("B200SD_ATTN_PAIR_BAR");
        cached = (e && e[0] == '0') ? 0 : 1;
    }
    return cached;
}

static int emu_mask_from_env() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("B200SD_ATTN_EXP_EMU");
        cached = (e && e[0] >= '0' && e[0] <= '4') ? e[0] - '0' : 2;  // default: a quarter of the exponentials (4: an eighth)
    }
    return cached;
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
    kp.stagger = stagger_from_env();
    kp.pair_bar = pair_bar_from_env();
    kp.causal = p->causal ? 1 : 0;
    B200SD_CHECK_ARG(!p->causal || p->sq == p->skv, "attention: causal needs sq == skv");
    const int rows_per_cta = ATT_BM * ATT_QT;
    kp.heads = p->heads;
    kp.n_qpairs = (p->sq + rows_per_cta - 1) / rows_per_cta;
    const long long items = (long long)kp.n_qpairs * p->heads * p->batch;
    B200SD_CHECK_ARG(items < (1LL << 31), "attention: too many work items");
    kp.items = (int)items;
    // Short kv loops are dominated by per-CTA set-up: there one CTA per SM walks the items.  Long ones (measured at 72 kv
    // tiles: -4 % as persistent CTAs, the static round-robin loses to the hardware's dynamic CTA scheduling) get a CTA each.
    const bool persistent = kp.n_kv_tiles <= 18;
    dim3 grid((unsigned)(persistent && items > num_sms() ? num_sms() : items), 1, 1);
    cudaStream_t s = as_stream(stream);
    // share of the exponentials evaluated on the FMA pipe instead of the MUFU: 0 none, 1 = 50 %, 2 = 25 %, 3 = 37.5 %
    const int emu = emu_mask_from_env();
    const bool quad = layout_from_env() != 0;
#define B200SD_ATTN_LAUNCH(F, M)                                                                               \
    return quad ? (persistent ? launch_attention<F, M, true, true>(tmQ, tmK, tmV, kp, grid, s)                  \
                              : launch_attention<F, M, false, true>(tmQ, tmK, tmV, kp, grid, s))                \
                : (persistent ? launch_attention<F, M, true, false>(tmQ, tmK, tmV, kp, grid, s)                 \
                              : launch_attention<F, M, false, false>(tmQ, tmK, tmV, kp, grid, s));
#define B200SD_ATTN_DISPATCH(F)                                                                              \
    switch (emu) {                                                                                            \
        case 1: B200SD_ATTN_LAUNCH(F, 0b10101010)                                                             \
        case 2: B200SD_ATTN_LAUNCH(F, 0b10001000)                                                             \
        case 3: B200SD_ATTN_LAUNCH(F, 0b10010010)                                                             \
        case 4: B200SD_ATTN_LAUNCH(F, 0b10000000)                                                             \
        default: B200SD_ATTN_LAUNCH(F, 0)                                                                     \
    }
    if (p->dtype == B200SD_F16) {
        B200SD_ATTN_DISPATCH(true)
    }
    B200SD_ATTN_DISPATCH(false)
#undef B200SD_ATTN_LAUNCH
#undef B200SD_ATTN_DISPATCH
}
