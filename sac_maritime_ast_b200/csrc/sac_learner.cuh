// SAC learner kernels + C ABI (sit_sac_*, include/shipsim.h).  Included at the end of shipsim.cu (one translation unit:
// shares fail()/CUDA_TRY, philox4x32_10 and the SitPolicy layout with the rollout kernels).
//
// Replaces the reference's `agent.update_parameters(memory, batch_size, updates)` (main_ast.py:350-357; the agent class
// lives in the absent ast_sac package, ast_core/algorithms/sac.py is empty) with standard twin-Q soft actor-critic on the
// device replay ring the rollout fills.  Policy network = the reference's Gaussian policy MLP (nn_models/mlp.py:95-148,
// distributions/normal.py:14-15,78-112, policies/gaussian_policy.py:65-148); hyper-parameters main_ast.py:41-68.
//
// Data layout.  Every dense layer is ONE row-major block [K+1][N]: K weight rows followed by the bias row, i.e. the weight
// matrix augmented with the bias.  Activations carry a trailing column of ones ([B][K+1]), so
//     forward   Y = relu(Xaug . Waug)            (bias folded into the product)
//     backward  dWaug = Xaug^T . dY              (weight and bias gradients in one product)
//               dX    = (dY . W^T) * (X > 0)     (first K rows of Waug only)
// are all instances of one strided, z-batched (twin critics side by side) GEMM kernel with a fused epilogue.
// An update is 28 such products + 9 element-wise kernels; sizes are tiny (B = 64, H = 256: 0.16 GFLOP), so the cost is
// launch latency: the whole update is captured once into a CUDA graph and replayed, with its independent parts as PARALLEL
// BRANCHES of that graph (two internal streams forked / joined by events during capture): the critics' forward on (s, a)
// beside the target computation, the policy's forward on s beside the critics' backward + Adam, and the three
// weight-gradient products of every backward pass beside its input-gradient chain (4 520 -> 6 340 updates/s at batch 64,
// 2 060 -> 2 570 at batch 4096).  All reductions have a fixed order: two runs with the same seed give bit-identical parameters.
#pragma once

namespace sac {

constexpr int S = SIT_N_OBS;       // observation width
constexpr int SA = SIT_N_OBS + 1;  // critic input width (one action, MSRL_env_ex.py:79-83)
constexpr float LOG_SIG_MIN = -20.0f, LOG_SIG_MAX = 2.0f;   // distributions/normal.py:14-15
constexpr float SQUASH_EPS = 1e-6f;                          // policies/gaussian_policy.py:16
constexpr float ADAM_B1 = 0.9f, ADAM_B2 = 0.999f, ADAM_EPS = 1e-8f;
constexpr float TARGET_ENTROPY = -1.0f;                      // -dim(action space)

// ---- the one GEMM: C[z](m, n) = epi( sum_k A[z](m, k) * B[z](k, n) ) --------------------------------------------------
// strides in elements; AK: A is contiguous along k (else along m); BN: B is contiguous along n (else along k) -- selects
// which thread index runs along the contiguous direction of the global loads.  32x32x32 tiles, 256 threads, 2x2 outputs
// per thread, fixed k order (deterministic).
enum { EPI_STORE = 0, EPI_RELU = 1, EPI_MASK = 2 };
struct Gemm {
    const float *A, *B; float* C; const float* Hm;
    int M, N, K;
    int a_rs, a_cs, b_rs, b_cs, ldc, ldh;
    long long a_zs, b_zs, c_zs, h_zs;
    int epi;
};

// The products are latency-bound (M = batch = 64, K <= 257, 16-32 CTAs): the k loop is a chain of dependent
// global-load -> barrier -> FMA rounds, so the tile is 128 deep -- 32 independent loads per thread in flight per round, three
// rounds for K = 257 instead of nine -- which took an H x H layer from ~20 us to a few us.  The k order of the sums is
// unchanged (ascending), so results are bit-identical to the 32-deep version.
constexpr int GEMM_BK = 128;
template <bool AK, bool BN>
__global__ void __launch_bounds__(256) gemm_kernel(const Gemm g) {
    __shared__ float As[GEMM_BK][33], Bs[GEMM_BK][33];   // [k][m], [k][n]
    const int t = threadIdx.x, z = blockIdx.z;
    const float* A = g.A + (long long)z * g.a_zs;
    const float* B = g.B + (long long)z * g.b_zs;
    const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
    const int ty = t >> 4, tx = t & 15;
    float c00 = 0.f, c01 = 0.f, c10 = 0.f, c11 = 0.f;
    const int lo = t & 31, hi = t >> 5;        // hi: 0..7
    for (int k0 = 0; k0 < g.K; k0 += GEMM_BK) {
        const int kt = min(GEMM_BK, g.K - k0);          // valid depth of this tile
        if (AK) {      // A contiguous along k: lanes run along k, 4 k-columns x 4 m-rows per thread
#pragma unroll
            for (int kc = 0; kc < GEMM_BK; kc += 32)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int ka = kc + lo, ma = hi + 8 * i, gm = m0 + ma;
                    As[ka][ma] = (gm < g.M && ka < kt) ? A[(long long)gm * g.a_rs + (long long)(k0 + ka) * g.a_cs] : 0.f;
                }
        } else {       // A contiguous along m: lanes run along m
#pragma unroll
            for (int i = 0; i < GEMM_BK / 8; ++i) {
                const int ka = hi + 8 * i, ma = lo, gm = m0 + ma;
                As[ka][ma] = (gm < g.M && ka < kt) ? A[(long long)gm * g.a_rs + (long long)(k0 + ka) * g.a_cs] : 0.f;
            }
        }
        if (BN) {      // B contiguous along n: lanes run along n
#pragma unroll
            for (int i = 0; i < GEMM_BK / 8; ++i) {
                const int kb = hi + 8 * i, nb = lo, gn = n0 + nb;
                Bs[kb][nb] = (gn < g.N && kb < kt) ? B[(long long)(k0 + kb) * g.b_rs + (long long)gn * g.b_cs] : 0.f;
            }
        } else {       // B contiguous along k
#pragma unroll
            for (int kc = 0; kc < GEMM_BK; kc += 32)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int kb = kc + lo, nb = hi + 8 * i, gn = n0 + nb;
                    Bs[kb][nb] = (gn < g.N && kb < kt) ? B[(long long)(k0 + kb) * g.b_rs + (long long)gn * g.b_cs] : 0.f;
                }
        }
        __syncthreads();
        const int kend = (kt + 31) & ~31;               // rows past kt are zero-filled above
        for (int kk0 = 0; kk0 < kend; kk0 += 32) {
#pragma unroll
            for (int kk = kk0; kk < kk0 + 32; ++kk) {
                const float a0 = As[kk][2 * ty], a1 = As[kk][2 * ty + 1], b0 = Bs[kk][2 * tx], b1 = Bs[kk][2 * tx + 1];
                c00 = fmaf(a0, b0, c00); c01 = fmaf(a0, b1, c01); c10 = fmaf(a1, b0, c10); c11 = fmaf(a1, b1, c11);
            }
        }
        __syncthreads();
    }
    float* C = g.C + (long long)z * g.c_zs;
    const float* Hm = g.epi == EPI_MASK ? g.Hm + (long long)z * g.h_zs : nullptr;
    const float acc[2][2] = {{c00, c01}, {c10, c11}};
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int m = m0 + 2 * ty + i, n = n0 + 2 * tx + j;
            if (m < g.M && n < g.N) {
                float v = acc[i][j];
                if (g.epi == EPI_RELU) v = fmaxf(v, 0.f);
                else if (g.epi == EPI_MASK) v = Hm[(long long)m * g.ldh + n] > 0.f ? v : 0.f;
                C[(long long)m * g.ldc + n] = v;
            }
        }
}

// ---- two degenerate shapes of the large-batch update, kept off the tiled kernels -------------------------------------
// k <= 16 (layer 0 forward: k = 11 / 12 inputs; input gradient through the 1- / 2-wide output layer): a CTA owns 32 output
// rows (their A values in shared memory), a thread one column n of all 8 (its B column in registers), lanes along n.  The
// sum runs over ascending k from zero with fmaf, exactly like gemm_kernel: bit-identical results.
constexpr int THIN_ROWS = 32;
__global__ void __launch_bounds__(256) gemm_thin_k_kernel(const Gemm g) {
    __shared__ float sa[THIN_ROWS][16];               // the CTA's rows of A (k <= 16)
    const int z = blockIdx.z, tid = threadIdx.x, m0 = blockIdx.x * THIN_ROWS;
    const float* A = g.A + (long long)z * g.a_zs;
    const float* B = g.B + (long long)z * g.b_zs;
    float* C = g.C + (long long)z * g.c_zs;
    const float* Hm = g.epi == EPI_MASK ? g.Hm + (long long)z * g.h_zs : nullptr;
    for (int j = tid; j < THIN_ROWS * 16; j += 256) {
        const int r = j >> 4, k = j & 15;
        sa[r][k] = (m0 + r < g.M && k < g.K) ? A[(long long)(m0 + r) * g.a_rs + (long long)k * g.a_cs] : 0.f;
    }
    __syncthreads();
    for (int n = tid; n < g.N; n += 256) {
        float b[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) b[k] = k < g.K ? B[(long long)k * g.b_rs + (long long)n * g.b_cs] : 0.f;
        for (int r0 = 0; r0 < THIN_ROWS; r0 += 8) {
            float hm[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) hm[r] = (Hm && m0 + r0 + r < g.M) ? Hm[(long long)(m0 + r0 + r) * g.ldh + n] : 1.f;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int m = m0 + r0 + r;
                if (m >= g.M) break;
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < 16; ++k) if (k < g.K) acc = fmaf(sa[r0 + r][k], b[k], acc);
                if (g.epi == EPI_RELU) acc = fmaxf(acc, 0.f);
                else if (g.epi == EPI_MASK) acc = hm[r] > 0.f ? acc : 0.f;
                C[(long long)m * g.ldc + n] = acc;
            }
        }
    }
}
// n <= 4 with A contiguous along k (the 1- / 2-wide output layer forward, d loss / d action): one warp per output row, lanes
// along k (coalesced), fixed shuffle tree -> deterministic; fp32 throughout.
__global__ void __launch_bounds__(256) gemm_skinny_n_kernel(const Gemm g) {
    const int z = blockIdx.z, lane = threadIdx.x & 31;
    const int m = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (m >= g.M) return;
    const float* A = g.A + (long long)z * g.a_zs + (long long)m * g.a_rs;
    const float* B = g.B + (long long)z * g.b_zs;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k = lane; k < g.K; k += 32) {
        const float a = A[k];
#pragma unroll
        for (int j = 0; j < 4; ++j) if (j < g.N) acc[j] = fmaf(a, B[(long long)k * g.b_rs + (long long)j * g.b_cs], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], d);
    if (lane < g.N) {
        float v = lane == 0 ? acc[0] : (lane == 1 ? acc[1] : (lane == 2 ? acc[2] : acc[3]));
        if (g.epi == EPI_RELU) v = fmaxf(v, 0.f);
        else if (g.epi == EPI_MASK) v = (g.Hm + (long long)z * g.h_zs)[(long long)m * g.ldh + lane] > 0.f ? v : 0.f;
        (g.C + (long long)z * g.c_zs)[(long long)m * g.ldc + lane] = v;
    }
}

// ---- the same GEMM on the tensor cores (tcgen05) for the large-batch learner -----------------------------------------
// C[z](m, n) = epi( sum_k A[z](m, k) * B[z](k, n) ) with bf16 operands and fp32 accumulation in TMEM.  Operands stay fp32 in
// global memory (master weights, activations, gradients): a CTA converts its tiles to bf16 on the way into shared memory and
// writes them in the UMMA canonical K-major layout (8-row x 16-byte core matrices, no swizzle), transposing where the source
// is contiguous along m / n instead of k -- so forward (X . W), input gradient (dY . W^T) and weight gradient (X^T . dY, the
// contraction over the batch) all run through this one kernel.  One CTA = one 128 x BN output tile (BN = 16 .. 128); the k
// range is cut into 64-deep chunks, double buffered: while tcgen05.mma (issued by one thread, completion signalled on an
// mbarrier by tcgen05.commit) consumes chunk c, all 256 threads stage chunk c + 1.  Products whose k is the batch are also
// split over `splits` CTAs along k: every CTA writes its fp32 partial tile and splitk_reduce_kernel sums the partials IN SPLIT
// ORDER and applies the epilogue -- a fixed summation order, so two runs give bit-identical results.
constexpr int TCG_M = 128, TCG_KC = 64, TCG_MAX_SPLITS = 32, TCG_MAX_STAGES = 5;
struct GemmTC { Gemm g; int bn, splits, tmem_cols, stages; float* partial; };

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT_TCG:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE_TCG;\n\t"
                 "bra LAB_WAIT_TCG;\n\tDONE_TCG:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}

__global__ void __launch_bounds__(256, 2) gemm_tc_kernel(const GemmTC t) {
    extern __shared__ __align__(128) unsigned char tcg_smem[];
    __shared__ uint64_t s_bar[TCG_MAX_STAGES];
    __shared__ uint32_t s_tmem;
    const Gemm& g = t.g;
    const int BN = t.bn, bn_shift = 31 - __clz(BN);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int z = blockIdx.z / t.splits, sp = blockIdx.z % t.splits;
    const int m0 = blockIdx.y * TCG_M, n0 = blockIdx.x * BN;
    const int n_chunks = (g.K + TCG_KC - 1) / TCG_KC, per = (n_chunks + t.splits - 1) / t.splits;
    const int c_lo = sp * per, c_hi = min(n_chunks, c_lo + per);
    const float* A = g.A + (long long)z * g.a_zs;
    const float* B = g.B + (long long)z * g.b_zs;
    const uint32_t a_bytes = TCG_M * TCG_KC * 2, stage_bytes = a_bytes + (uint32_t)BN * TCG_KC * 2;
    float* s_c = reinterpret_cast<float*>(tcg_smem);              // epilogue tile [128][BN + 1], aliases the staging buffers

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(t.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 32) {
        for (int i = 0; i < TCG_MAX_STAGES; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    const uint32_t tmem_base = s_tmem;
    // kind::f16 instruction descriptor: fp32 accumulate, bf16 A and B, both K-major, N = BN, M = 128
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TCG_M >> 4) << 24);
    const uint32_t lbo = 128u, sbo = 8u * 128u;                   // next 16-byte k chunk | next 8-row group (64-deep chunk)

    // One 64-deep chunk of an operand tile (`rows` = 128 for A, BN for B) goes global -> registers -> canonical K-major bf16
    // image in two phases, so that ALL of a chunk's loads are in flight together and the loads of chunk c + 1 are issued
    // before the CTA waits for the tensor core to release chunk c - 1's buffer.  Element (r, k) = P[(r0 + r) rs + (k0 + k) cs].
    //   mode 1 (source contiguous along k, 16-byte aligned rows): a thread owns (r, 8 consecutive k): two 128-bit loads;
    //   mode 2 (source contiguous along r, 16-byte aligned):      a thread owns (4 consecutive r, 8 k): eight 128-bit loads,
    //                                                             transposed in registers into four core-matrix rows;
    //   mode 0 (anything else, and the k tail of mode 1):         a thread owns (r, 8 k): eight predicated scalar loads.
    struct Regs { float v[32]; };
    auto pick_mode = [&](const float* P, const long long rs, const long long cs) {
        const bool al = (reinterpret_cast<uintptr_t>(P) & 15) == 0;
        if (cs == 1 && al && (rs & 3) == 0) return 1;
        if (rs == 1 && al && (cs & 3) == 0) return 2;
        return 0;
    };
    auto load_tile = [&](Regs& R, int mode, const float* P, const int r_total, const long long rs, const long long cs, const int r0,
                         const int rows_shift, const int k0) {
        const int rows = 1 << rows_shift;
        if (mode == 1 && k0 + TCG_KC > g.K) mode = 0;
        if (mode == 2) {
            const int q_shift = rows_shift - 2;                       // row quads per tile
            const int r4 = (tid & ((1 << q_shift) - 1)) * 4, ko = tid >> q_shift, gr = r0 + r4;
            const bool ok = tid < (8 << q_shift);
            const float* p = P + gr + (long long)(k0 + ko * 8) * cs;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ok && gr < r_total && k0 + ko * 8 + j < g.K) q = *reinterpret_cast<const float4*>(p + (long long)j * cs);
                R.v[j * 4 + 0] = q.x;
                R.v[j * 4 + 1] = gr + 1 < r_total ? q.y : 0.f;
                R.v[j * 4 + 2] = gr + 2 < r_total ? q.z : 0.f;
                R.v[j * 4 + 3] = gr + 3 < r_total ? q.w : 0.f;
            }
            return;
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int item = tid + it * 256;
            const int r = item & (rows - 1), ko = item >> rows_shift, gr = r0 + r, kb = k0 + ko * 8;
            const bool ok = item < 8 * rows && gr < r_total;
            const float* p = P + (long long)gr * rs + (long long)kb * cs;
            if (mode == 1) {
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
                if (ok) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 4); }
                R.v[it * 8 + 0] = a.x; R.v[it * 8 + 1] = a.y; R.v[it * 8 + 2] = a.z; R.v[it * 8 + 3] = a.w;
                R.v[it * 8 + 4] = b.x; R.v[it * 8 + 5] = b.y; R.v[it * 8 + 6] = b.z; R.v[it * 8 + 7] = b.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) R.v[it * 8 + j] = (ok && kb + j < g.K) ? p[(long long)j * cs] : 0.f;
            }
        }
    };
    auto store_tile = [&](const Regs& R, const int mode, const int rows_shift, unsigned char* dst) {
        const int rows = 1 << rows_shift;
        if (mode == 2) {
            const int q_shift = rows_shift - 2;
            const int r4 = (tid & ((1 << q_shift) - 1)) * 4, ko = tid >> q_shift;
            if (tid < (8 << q_shift)) {
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) {
                    const int r = r4 + rr;
                    uint4 w;
                    w.x = pack_bf16x2(R.v[0 * 4 + rr], R.v[1 * 4 + rr]); w.y = pack_bf16x2(R.v[2 * 4 + rr], R.v[3 * 4 + rr]);
                    w.z = pack_bf16x2(R.v[4 * 4 + rr], R.v[5 * 4 + rr]); w.w = pack_bf16x2(R.v[6 * 4 + rr], R.v[7 * 4 + rr]);
                    *reinterpret_cast<uint4*>(dst + (size_t)(r >> 3) * sbo + (size_t)ko * lbo + (r & 7) * 16) = w;
                }
            }
            return;
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int item = tid + it * 256;
            if (item < 8 * rows) {
                const int r = item & (rows - 1), ko = item >> rows_shift;
                const float* v = &R.v[it * 8];
                uint4 w;
                w.x = pack_bf16x2(v[0], v[1]); w.y = pack_bf16x2(v[2], v[3]); w.z = pack_bf16x2(v[4], v[5]); w.w = pack_bf16x2(v[6], v[7]);
                *reinterpret_cast<uint4*>(dst + (size_t)(r >> 3) * sbo + (size_t)ko * lbo + (r & 7) * 16) = w;
            }
        }
    };
    const int mode_a = pick_mode(A, g.a_rs, g.a_cs), mode_b = pick_mode(B, g.b_cs, g.b_rs);

    // `stages` buffers (2 .. 5): a CTA whose k range fits them never waits for the tensor core inside the loop -- the issue ->
    // completion latency of a tcgen05.mma batch is far longer than its math at these tile sizes
    uint32_t phase[TCG_MAX_STAGES] = {0u, 0u, 0u, 0u, 0u};
    bool used[TCG_MAX_STAGES] = {false, false, false, false, false};
    int last_b = 0;
    Regs ra, rb;
    if (c_lo < c_hi) {
        load_tile(ra, mode_a, A, g.M, g.a_rs, g.a_cs, m0, 7, c_lo * TCG_KC);
        load_tile(rb, mode_b, B, g.N, g.b_cs, g.b_rs, n0, bn_shift, c_lo * TCG_KC);
    }
    for (int c = c_lo; c < c_hi; ++c) {
        const int b = (c - c_lo) % t.stages;
        unsigned char* sA = tcg_smem + (size_t)b * stage_bytes;
        unsigned char* sB = sA + a_bytes;
        if (used[b]) { mbar_wait(smem_u32(&s_bar[b]), phase[b]); phase[b] ^= 1u; }     // the MMAs that read this buffer are done
        store_tile(ra, mode_a, 7, sA);
        store_tile(rb, mode_b, bn_shift, sB);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::);
            const uint64_t da = umma_desc_kmajor_noswizzle(smem_u32(sA), lbo, sbo), db = umma_desc_kmajor_noswizzle(smem_u32(sB), lbo, sbo);
#pragma unroll
            for (int ks = 0; ks < TCG_KC / 16; ++ks) {
                const uint64_t adv = (uint64_t)((ks * 2 * lbo) >> 4);
                const uint32_t acc = (c > c_lo || ks > 0) ? 1u : 0u;
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             ::"r"(tmem_base), "l"(da + adv), "l"(db + adv), "r"(idesc), "r"(acc) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar[b])) : "memory");
        }
        used[b] = true; last_b = b;
        if (c + 1 < c_hi) {          // next chunk's loads fly while the tensor core works on this one
            load_tile(ra, mode_a, A, g.M, g.a_rs, g.a_cs, m0, 7, (c + 1) * TCG_KC);
            load_tile(rb, mode_b, B, g.N, g.b_cs, g.b_rs, n0, bn_shift, (c + 1) * TCG_KC);
        }
    }
    const bool any = c_hi > c_lo;
    if (any) mbar_wait(smem_u32(&s_bar[last_b]), phase[last_b]);   // commits complete in order: every MMA of this CTA is done
    asm volatile("tcgen05.fence::after_thread_sync;" ::);

    // accumulator -> shared tile (thread = TMEM lane = output row; the two thread halves take the two column halves)
    {
        const int row = tid & (TCG_M - 1), half = tid >> 7, ch = BN >> 1;
        const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        for (int c0 = half * ch; c0 < (half + 1) * ch; c0 += 8) {
            uint32_t r[8];
            if (any) {
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                             : "r"(taddr + (uint32_t)c0));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) r[j] = 0u;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) s_c[row * (BN + 1) + c0 + j] = __uint_as_float(r[j]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();

    if (t.splits == 1) {
        float* C = g.C + (long long)z * g.c_zs;
        const float* Hm = g.epi == EPI_MASK ? g.Hm + (long long)z * g.h_zs : nullptr;
        // coalesced stores, eight outputs per thread and trip so that the mask loads of a trip are in flight together
        for (int i0 = tid; i0 < TCG_M * BN; i0 += 256 * 8) {
            float hm[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = i0 + u * 256, r = idx >> bn_shift, cc = idx & (BN - 1), m = m0 + r, n = n0 + cc;
                hm[u] = (g.epi == EPI_MASK && idx < TCG_M * BN && m < g.M && n < g.N) ? Hm[(long long)m * g.ldh + n] : 1.f;
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = i0 + u * 256, r = idx >> bn_shift, cc = idx & (BN - 1), m = m0 + r, n = n0 + cc;
                if (idx < TCG_M * BN && m < g.M && n < g.N) {
                    float v = s_c[r * (BN + 1) + cc];
                    if (g.epi == EPI_RELU) v = fmaxf(v, 0.f);
                    else if (g.epi == EPI_MASK) v = hm[u] > 0.f ? v : 0.f;
                    C[(long long)m * g.ldc + n] = v;
                }
            }
        }
    } else {
        // split-k: this CTA's partial tile -> [z][split][M][N]; splitk_reduce_kernel sums the splits in order
        float* P = t.partial + ((long long)(z * t.splits + sp) * g.M) * g.N;
        for (int idx = tid; idx < TCG_M * BN; idx += 256) {
            const int r = idx >> bn_shift, cc = idx & (BN - 1), m = m0 + r, n = n0 + cc;
            if (m < g.M && n < g.N) P[(long long)m * g.N + n] = s_c[r * (BN + 1) + cc];
        }
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(t.tmem_cols));
}

// C[z](m, n) = epi( sum over splits, IN SPLIT ORDER, of the partial tiles of gemm_tc_kernel ): one thread per output, all of
// its loads in flight together; a fixed summation order, so two runs give bit-identical results.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const GemmTC t) {
    const Gemm& g = t.g;
    const long long mn = (long long)g.M * g.N;
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    const int z = blockIdx.y;
    if (i >= mn) return;
    const float* pp = t.partial + (long long)z * t.splits * mn + i;
    float acc = 0.f;
    for (int q0 = 0; q0 < t.splits; q0 += 8) {
        float x[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) x[q] = q0 + q < t.splits ? pp[(q0 + q) * mn] : 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) acc += x[q];
    }
    const int m = (int)(i / g.N), n = (int)(i % g.N);
    if (g.epi == EPI_RELU) acc = fmaxf(acc, 0.f);
    else if (g.epi == EPI_MASK) acc = (g.Hm + (long long)z * g.h_zs)[(long long)m * g.ldh + n] > 0.f ? acc : 0.f;
    (g.C + (long long)z * g.c_zs)[(long long)m * g.ldc + n] = acc;
}

// ---- fused forward of one 3-layer network on the tensor-core path ------------------------------------------------------
// out = W2^T relu(W1^T relu(W0^T x)) for 128 batch rows per CTA and `Z` networks (grid.y), the hidden activations h1, h2
// written to HBM on the way for the backward pass.  One kernel instead of three (thin-k, tcgen05 tile GEMM, skinny-n):
//   layer 0 (k = 11 / 12 inputs) on the CUDA cores, fp32, ascending k like gemm_thin_k_kernel; the relu'd row goes to h1
//     (fp32) and, as bf16, into the canonical K-major image of the A operand in shared memory;
//   layer 1 (H x H) on tcgen05: the bf16 image of W1 (pack_w1_nets_kernel, canonical K-major, rows = output unit) lands with
//     ONE bulk async copy while layer 0 is computed; 128 x H fp32 accumulator in TMEM; the bias row of the augmented layer
//     is added in fp32 in the epilogue;
//   layer 2 (1 or 2 outputs) folded into the epilogue: the thread that owns a row reads its accumulator columns, applies
//     bias + relu, writes h2 and accumulates the outputs in fp32, ascending k inside each column half, the two halves added
//     in a fixed order -- deterministic.
struct FwdTC {
    const float* X; int kin1;                 // [B][kin1] input rows, last column = 1 (the bias input)
    const float* W; long long w_zs;           // network block (augmented layers at l0, l1, l2), stride between the Z networks
    long long l0, l1, l2;
    const unsigned char* w1p; long long w1p_zs;   // bf16 image of W1's H x H part, bytes
    float *H1, *H2, *Out; long long hz;       // activations [Z][B][ld] (nullptr: not kept -- a pass no backward goes through), outputs [Z][B][nout]
    unsigned char* IMG1; long long img_zs;    // bf16 tile image of h1 for d W1 (nullptr: none), bytes between the Z networks
    unsigned char* IMG0;                      // ... of the input rows for d W0 (one 128-feature tile, shared by the Z networks)
    uint32_t *M1, *M2;                        // sign bits of h1, h2 for the fused backward: [Z][B][8 words], bit k of a row = word k >> 5, bit k & 31; or nullptr
    int B, H, ld, nout, tmem_cols;
    // policy only (nullptr otherwise): the tanh-Gaussian head of head_kernel in the epilogue -- a = tanh(mu + sigma eps) into the
    // action column of a critic input buffer [B][SA + 1], log pi into logp [B]
    const float* eps; float* xa; float* logp;
};
// bf16 K-major images of the H x H part W1[k][n] of layer 1 (entry q: weights at ws + w_off[q], image number img[q] of H H 2
// bytes each).  Forward image (by_k = 0): operand rows = output unit n, contraction along k.  Backward image (by_k = 1):
// operand rows = input unit k, contraction along n (d h1 = d h2 . W1^T).
#define SAC_W1_IMAGES 8   /* forward: policy, q1, q2, q1 target, q2 target; backward: policy, q1, q2 */
struct PackNets { long long w_off[SAC_W1_IMAGES]; int img[SAC_W1_IMAGES], by_k[SAC_W1_IMAGES]; };
__global__ void __launch_bounds__(256) pack_w1_nets_kernel(const float* ws, unsigned char* w1p, const PackNets pn, const int H) {
    const float* w1 = ws + pn.w_off[blockIdx.y];
    unsigned char* dst = w1p + (size_t)pn.img[blockIdx.y] * H * H * 2;
    const uint32_t sbo = 16u * (uint32_t)H, lbo = 128u;
    const bool by_k = pn.by_k[blockIdx.y] != 0;
    for (int item = blockIdx.x * 256 + threadIdx.x; item < H * (H >> 3); item += gridDim.x * 256) {
        uint4 v;
        int r, c;                                   // operand row, 8-wide chunk along the contraction
        if (!by_k) {                                // row = n, chunk of 8 consecutive k (stride H in memory)
            r = item % H; c = item / H;
            const float* w = w1 + (size_t)(c * 8) * H + r;
            v.x = pack_bf16x2(w[0], w[H]); v.y = pack_bf16x2(w[2 * H], w[3 * H]);
            v.z = pack_bf16x2(w[4 * H], w[5 * H]); v.w = pack_bf16x2(w[6 * H], w[7 * H]);
        } else {                                    // row = k, chunk of 8 consecutive n (contiguous in memory)
            c = item % (H >> 3); r = item / (H >> 3);
            const float* w = w1 + (size_t)r * H + c * 8;
            v.x = pack_bf16x2(w[0], w[1]); v.y = pack_bf16x2(w[2], w[3]); v.z = pack_bf16x2(w[4], w[5]); v.w = pack_bf16x2(w[6], w[7]);
        }
        *reinterpret_cast<uint4*>(dst + (size_t)(r >> 3) * sbo + (size_t)c * lbo + (r & 7) * 16) = v;
    }
}

// Row masks: sign bits of a row's H activations as 8 words, bit k of the row = word k >> 5, bit k & 31.
__device__ __forceinline__ uint32_t mask_word(const uint4& lo, const uint4& hi, const int word) {
    return word < 4 ? (word < 2 ? (word == 0 ? lo.x : lo.y) : (word == 2 ? lo.z : lo.w))
                    : (word < 6 ? (word == 4 ? hi.x : hi.y) : (word == 6 ? hi.z : hi.w));
}
// Epilogue stores: a thread owns one accumulator row (TMEM lane), so its 16 values of a chunk would go out as 16-byte pieces
// of 32 different rows per instruction.  The warp transposes the chunk through a private [32][16] staging tile (row stride
// 20 floats: conflict-free 128-bit accesses) and stores 8 rows x 64 contiguous bytes per instruction instead.
#define FUSED_STG_STRIDE 20
// CTA shape of the fused kernels: 16 warps (the 8-warp version issued 32 % of the time: two warps per scheduler cannot hide
// the dependent-issue latency of the layer-0 FMA chains and the epilogue); a warp owns 8 rows in the layer-0 / layer-2 phase,
// the four thread quarters take the four column quarters of the accumulator in the epilogue
constexpr int FUSED_NT = 512, FUSED_NW = FUSED_NT / 32, FUSED_RB = TCG_M / (8 * FUSED_NW), FUSED_NQ = FUSED_NT / TCG_M;
// img != nullptr: the 32 x 16 tile also goes, as bf16, into an operand image (see image_store_warp): block base of feature tile
// f0 / 128 and of the tile's 64-row chunk already applied by the caller, f0 = first feature, r0 = first row inside the chunk
__device__ __forceinline__ void staged_store16(float* stg, const int lane, const float (&v)[16], float* gbase, const long long ld,
                                               const int rows_left, unsigned char* img = nullptr, const int f0 = 0, const int r0 = 0) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(stg + lane * FUSED_STG_STRIDE + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    __syncwarp();
    if (img) {      // lane = (feature of the 16, one of two 8-row groups per pass): 8 rows of one feature = one 16-byte piece
        const int c = lane & 15, g = lane >> 4, f = (f0 & 127) + c;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const int rg = 2 * p + g;
            const float* col = stg + (rg * 8) * FUSED_STG_STRIDE + c;
            uint4 w;
            w.x = pack_bf16x2(col[0], col[FUSED_STG_STRIDE]); w.y = pack_bf16x2(col[2 * FUSED_STG_STRIDE], col[3 * FUSED_STG_STRIDE]);
            w.z = pack_bf16x2(col[4 * FUSED_STG_STRIDE], col[5 * FUSED_STG_STRIDE]); w.w = pack_bf16x2(col[6 * FUSED_STG_STRIDE], col[7 * FUSED_STG_STRIDE]);
            *reinterpret_cast<uint4*>(img + (size_t)(f >> 3) * 1024 + (size_t)((r0 >> 3) + rg) * 128 + (f & 7) * 16) = w;
        }
    }
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
        const int rr = pass * 8 + (lane >> 2), cc = (lane & 3) * 4;
        const float4 q = *reinterpret_cast<const float4*>(stg + rr * FUSED_STG_STRIDE + cc);
        if (gbase && rr < rows_left) *reinterpret_cast<float4*>(gbase + (long long)rr * ld + cc) = q;
    }
    __syncwarp();
}

// d W1 operand images: element (feature f, batch row m) of a network's [rows][H] matrix lives in block (f / 128, m / 64) of
// 16 KB at (f % 128 / 8) 1024 + (m % 64 / 8) 128 + (f % 8) 16 + (m % 8) 2 -- the canonical K-major core-matrix order with the
// batch as K (LBO 128, SBO 1024).  In the layer-0 / layer-2 phase a warp has just written four 8 x 8 core matrices
// (8 rows x 8 features, features contiguous) of the A image in shared memory; ldmatrix.trans hands every lane two batch rows
// of one feature per matrix, i.e. 4 contiguous bytes of the transposed order: four 32-bit stores per lane.
// `core` = shared address of the lane's own 16-byte row of its core matrix (lane = (row of the block, matrix)).
__device__ __forceinline__ void image_store_warp(unsigned char* img, const int n_chunks, const int k0_warp, const int m_blk, const uint32_t core,
                                                 const int lane) {
    uint32_t q0, q1, q2, q3;
    __syncwarp();
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(core));
    // matrix j = features k0_warp + 8 j .. + 7; this lane: feature lane / 4 of each, batch rows m_blk + 2 (lane % 4), + 1
    const int fr = lane >> 2, mr = m_blk + 2 * (lane & 3);
    const uint32_t q[4] = {q0, q1, q2, q3};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int f = k0_warp + 8 * j + fr;
        unsigned char* p = img + ((size_t)(f >> 7) * n_chunks + (mr >> 6)) * 16384 + (size_t)((f & 127) >> 3) * 1024 + (size_t)((mr & 63) >> 3) * 128 +
                           (f & 7) * 16 + (mr & 7) * 2;
        *reinterpret_cast<uint32_t*>(p) = q[j];
    }
}

__global__ void __launch_bounds__(FUSED_NT, 1) mlp_fwd_tc_kernel(const FwdTC f) {
    extern __shared__ __align__(128) unsigned char fw_smem[];
    const int H = f.H, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, z = blockIdx.y;
    const int m0 = blockIdx.x * TCG_M;
    const float* W = f.W + (long long)z * f.w_zs;
    unsigned char* sA = fw_smem;                                      // [128/8][H/8][8][8] bf16; after the MMAs: staging tiles
    unsigned char* sB = sA + (size_t)TCG_M * H * 2;                   // [H/8][H/8][8][8] bf16, rows = output unit
    float* s_w2 = reinterpret_cast<float*>(sB + (size_t)H * H * 2);   // [H + 1][2]
    float* s_b1 = s_w2 + 2 * (H + 1);                                 // [H]
    float* s_part = s_b1 + H;                                         // [FUSED_NQ - 1][128][2]
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_part + 2 * TCG_M * (FUSED_NQ - 1) + 2);   // [0]: MMAs done, [1]: W1 image landed (8-byte aligned)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 2);
    float* s_w0 = reinterpret_cast<float*>(s_bar + 4);                // [kin1][H] (the network blocks are only 4-byte aligned in the workspace)
    const uint32_t sbo = 16u * (uint32_t)H, lbo = 128u;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(f.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 32) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar + 1)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::);
        const uint32_t bytes = (uint32_t)H * (uint32_t)H * 2u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(s_bar + 1)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sB)), "l"(f.w1p + (long long)z * f.w1p_zs), "r"(bytes), "r"(smem_u32(s_bar + 1)) : "memory");
    }
    // layer 0 mapping: a warp owns FUSED_RB blocks of 8 rows; lane = (row of the block, 8-column group of a 32-column
    // step), so one instruction writes 8 rows x 128 contiguous bytes of h1 and 512 contiguous bytes of the A image
    const int rsub = lane & 7, grp = lane >> 3;
    float x[FUSED_RB][12];
#pragma unroll
    for (int rb = 0; rb < FUSED_RB; ++rb) {
        const int m = m0 + warp * (8 * FUSED_RB) + rb * 8 + rsub;
#pragma unroll
        for (int c = 0; c < 12; ++c) x[rb][c] = (m < f.B && c < f.kin1) ? f.X[(long long)m * f.kin1 + c] : 0.f;
        if (f.IMG0 && grp == 0 && z == 0) {      // the input rows as the A operand of d W0 (rows past the batch: zeros)
            unsigned char* p = f.IMG0 + (size_t)(m >> 6) * 16384 + (size_t)((m & 63) >> 3) * 128 + (m & 7) * 2;
#pragma unroll
            for (int c = 0; c < 12; ++c) {
                if (c < f.kin1) {
                    unsigned short hb;
                    asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(hb) : "f"(x[rb][c]));
                    *reinterpret_cast<unsigned short*>(p + (size_t)(c >> 3) * 1024 + (c & 7) * 16) = hb;
                }
            }
        }
    }
    {   // small operands -> shared memory with ALL loads of a thread in flight together (a loop of dependent
        // load -> store trips cost one L2 round trip each: 15 of them were 9 of this kernel's 19 us)
        constexpr int Q0 = 12 * 256 / FUSED_NT, Q2 = (2 * 257 + FUSED_NT - 1) / FUSED_NT;
        float t0[Q0], t2[Q2], t1 = 0.f;
        const int n0 = f.kin1 * H, n2 = (H + 1) * f.nout;              // <= 12 x 256 | <= 257 x 2
#pragma unroll
        for (int q = 0; q < Q0; ++q) { const int i = tid + q * FUSED_NT; t0[q] = i < n0 ? W[f.l0 + i] : 0.f; }
#pragma unroll
        for (int q = 0; q < Q2; ++q) { const int i = tid + q * FUSED_NT; t2[q] = i < n2 ? W[f.l2 + i] : 0.f; }
        if (tid < H) t1 = W[f.l1 + (long long)H * H + tid];
#pragma unroll
        for (int q = 0; q < Q0; ++q) { const int i = tid + q * FUSED_NT; if (i < n0) s_w0[i] = t0[q]; }
#pragma unroll
        for (int q = 0; q < Q2; ++q) {
            const int i = tid + q * FUSED_NT;
            if (i < n2) {
                const int k = i / f.nout, j = i % f.nout;
                s_w2[2 * k + j] = t2[q];
                if (f.nout == 1) s_w2[2 * k + 1] = 0.f;
            }
        }
        if (tid < H) s_b1[tid] = t1;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    const uint32_t tmem_base = *s_tmem;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(H >> 3) << 17) | ((uint32_t)(TCG_M >> 4) << 24);

    // layer 0: relu(x . W0), fp32, ascending input index
#pragma unroll
    for (int rb = 0; rb < FUSED_RB; ++rb) {
        const int row = warp * (8 * FUSED_RB) + rb * 8 + rsub, m = m0 + row;
        const bool live = m < f.B;
        float* h1 = f.H1 + (long long)z * f.hz + (long long)m * f.ld;
        for (int it = 0; it < (H >> 5); ++it) {
            const int k0 = it * 32 + grp * 8;
            float a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = 0.f;
#pragma unroll
            for (int c = 0; c < 12; ++c) {
                if (c < f.kin1) {
                    const float4 wa = *reinterpret_cast<const float4*>(s_w0 + c * H + k0);
                    const float4 wb = *reinterpret_cast<const float4*>(s_w0 + c * H + k0 + 4);
                    const float xc = x[rb][c];
                    a[0] = fmaf(xc, wa.x, a[0]); a[1] = fmaf(xc, wa.y, a[1]); a[2] = fmaf(xc, wa.z, a[2]); a[3] = fmaf(xc, wa.w, a[3]);
                    a[4] = fmaf(xc, wb.x, a[4]); a[5] = fmaf(xc, wb.y, a[5]); a[6] = fmaf(xc, wb.z, a[6]); a[7] = fmaf(xc, wb.w, a[7]);
                }
            }
            uint32_t bits = 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) { a[j] = fmaxf(a[j], 0.f); bits |= (a[j] > 0.f ? 1u : 0u) << j; }
            bits <<= 8 * grp;                                          // the four column groups of a row make up mask word `it`
            bits |= __shfl_xor_sync(0xffffffffu, bits, 8);
            bits |= __shfl_xor_sync(0xffffffffu, bits, 16);
            if (live) {                                                // (no backward through this pass: nothing is kept)
                if (grp == 0 && f.M1) f.M1[((long long)z * f.B + m) * 8 + it] = bits;
                if (f.H1 && !f.IMG1) {                                 // (with the operand image nobody reads the fp32 h1)
                    *reinterpret_cast<float4*>(h1 + k0) = make_float4(a[0], a[1], a[2], a[3]);
                    *reinterpret_cast<float4*>(h1 + k0 + 4) = make_float4(a[4], a[5], a[6], a[7]);
                }
            }
            uint4 v;
            v.x = pack_bf16x2(a[0], a[1]); v.y = pack_bf16x2(a[2], a[3]); v.z = pack_bf16x2(a[4], a[5]); v.w = pack_bf16x2(a[6], a[7]);
            unsigned char* core = sA + (size_t)(row >> 3) * sbo + (size_t)(k0 >> 3) * lbo + (row & 7) * 16;
            *reinterpret_cast<uint4*>(core) = v;
            // (rows past the batch are zero rows of the image: x = 0 there)
            if (f.IMG1) image_store_warp(f.IMG1 + (long long)z * f.img_zs, (int)gridDim.x * 2, it * 32, m - rsub, smem_u32(core), lane);
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    // layer 1 on the tensor cores: D[128 x H] (TMEM, fp32) = A[128 x H] . B[H x H]
    if (tid == 0) {
        mbar_wait(smem_u32(s_bar + 1), 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        const uint64_t da = umma_desc_kmajor_noswizzle(smem_u32(sA), lbo, sbo), db = umma_desc_kmajor_noswizzle(smem_u32(sB), lbo, sbo);
        for (int ks = 0; ks < (H >> 4); ++ks) {
            const uint64_t adv = (uint64_t)((ks * 2 * lbo) >> 4);
            const uint32_t acc = ks > 0 ? 1u : 0u;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tmem_base), "l"(da + adv), "l"(db + adv), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(s_bar)) : "memory");
    }
    mbar_wait(smem_u32(s_bar), 0u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    // epilogue: thread = accumulator row, the FUSED_NQ thread quarters take the column quarters; h2 = relu(D + b1) -> HBM through
    // the warp's staging tile (the A image is dead), output layer accumulated on the way
    const int row = tid & (TCG_M - 1), half = tid >> 7, kh = H / FUSED_NQ;      // (`half`: this thread's column part)
    const int m = m0 + row;
    const bool live = m < f.B;
    float* stg = reinterpret_cast<float*>(sA) + warp * 32 * FUSED_STG_STRIDE;
    const int wrow0 = m0 + (warp & 3) * 32;                            // first row of this warp's 32
    float* h2w = f.H2 + (long long)z * f.hz + (long long)wrow0 * f.ld;
    float o0 = 0.f, o1 = 0.f;
    uint32_t mword = 0u;
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (int c0 = half * kh; c0 < (half + 1) * kh; c0 += 16) {
        uint32_t r[16];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                       "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(taddr + (uint32_t)c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        float hv[16];
        uint32_t bits = 0u;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            hv[j] = fmaxf(__uint_as_float(r[j]) + s_b1[c0 + j], 0.0f);
            bits |= (hv[j] > 0.f ? 1u : 0u) << j;
            const float2 w = reinterpret_cast<const float2*>(s_w2)[c0 + j];
            o0 = fmaf(hv[j], w.x, o0);
            o1 = fmaf(hv[j], w.y, o1);
        }
        mword |= bits << (c0 & 31);                                    // kh is a multiple of 32: a word belongs to one column part
        if ((c0 & 16) && live && f.M2) { f.M2[((long long)z * f.B + m) * 8 + (c0 >> 5)] = mword; }
        if (c0 & 16) mword = 0u;
        if (f.H2) staged_store16(stg, lane, hv, h2w + c0, f.ld, f.B - wrow0);
    }
    if (half > 0) { s_part[2 * ((half - 1) * TCG_M + row)] = o0; s_part[2 * ((half - 1) * TCG_M + row) + 1] = o1; }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    if (half == 0 && live) {
        float* out = f.Out + (long long)z * f.B * f.nout + (long long)m * f.nout;
#pragma unroll
        for (int q = 0; q < FUSED_NQ - 1; ++q) { o0 += s_part[2 * (q * TCG_M + row)]; o1 += s_part[2 * (q * TCG_M + row) + 1]; }   // fixed order
        const float mu = o0 + s_w2[2 * H];
        out[0] = mu;
        if (f.nout == 2) {
            const float lsr = o1 + s_w2[2 * H + 1];
            out[1] = lsr;
            if (f.eps) {                                               // same arithmetic as head_kernel
                const float ls = fminf(fmaxf(lsr, LOG_SIG_MIN), LOG_SIG_MAX);
                const float eps = f.eps[m];
                const float a = tanhf(fmaf(expf(ls), eps, mu));
                f.xa[(long long)m * (SA + 1) + S] = a;
                f.logp[m] = -0.5f * eps * eps - ls - 0.91893853320467274f - logf(1.0f - a * a + SQUASH_EPS);
            }
        }
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(f.tmem_cols));
}

// ---- fused input-gradient chain of one 3-layer network on the tensor-core path ------------------------------------------
// d h2 = (d out . W2^T) masked by h2 > 0 (CUDA cores, ascending output index like gemm_thin_k_kernel) -> HBM (for the
// weight-gradient product of layer 1) and, as bf16, the A image;  d h1 = (d h2 . W1^T) masked by h1 > 0 on tcgen05 (W1's
// by-k bf16 image, one bulk copy) -> HBM (for layer 0's weight gradient);  optionally d x of ONE input column (the action
// of the critics: d Q / d a) accumulated in the epilogue, ascending k inside each column half, halves added in a fixed order.
// One kernel instead of thin-k + tile GEMM (+ skinny-n); the weight-gradient products stay on gemm_tc_kernel.
struct BwdTC {
    const float* DOut; int nout;              // [Z][B][nout]
    // seed != 0 (critics, nout = 1): d out computed here with the arithmetic of critic_loss_kernel (1) / policy_seed_kernel (2)
    // instead of read, so those kernels -- still run for the losses they report and for layer 2's weight gradient -- leave
    // the critical path
    int seed; const float *oq, *oqt, *logp_next, *msk, *rew, *alpha; float gamma;
    const float* W; long long w_zs; long long l0, l2;
    const unsigned char* w1p; long long w1p_zs;   // by-k bf16 image of W1
    const uint32_t *M1, *M2;                  // sign bits of the forward activations h1, h2 (mlp_fwd_tc_kernel): [Z][B][8 words]
    unsigned char* IMG2; long long img_zs;    // bf16 tile image of d h2 for d W1 (nullptr: none)
    unsigned char* IMG1;                      // ... of d h1 for d W0 (with IMG2)
    float* colsum;                            // [Z][row tiles][H] column sums of this CTA's d h2 rows (with IMG2)
    float *DH2, *DH1;                         // [Z][B][H], nullptr: not stored (no weight gradients wanted)
    float* DX; int dx_row;                    // [Z][B] or nullptr; row of W0 (input column) the gradient is taken for
    int B, H, tmem_cols;
};
__global__ void __launch_bounds__(FUSED_NT, 1) mlp_bwd_tc_kernel(const BwdTC f) {
    extern __shared__ __align__(128) unsigned char bw_smem[];
    const int H = f.H, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, z = blockIdx.y;
    const int m0 = blockIdx.x * TCG_M;
    const float* W = f.W + (long long)z * f.w_zs;
    unsigned char* sA = bw_smem;                                      // A image; after the MMAs: staging tiles
    unsigned char* sB = sA + (size_t)TCG_M * H * 2;
    float* s_w2 = reinterpret_cast<float*>(sB + (size_t)H * H * 2);   // [H][2]
    float* s_wx = s_w2 + 2 * H;                                       // [H]: row dx_row of W0
    float* s_part = s_wx + H;                                         // [FUSED_NQ - 1][128]
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_part + TCG_M * (FUSED_NQ - 1));
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 2);
    float* s_cs = reinterpret_cast<float*>(s_bar + 4);                // [FUSED_NW x FUSED_RB][H] column sums of 8-row blocks (IMG2 only)
    const uint32_t sbo = 16u * (uint32_t)H, lbo = 128u;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(f.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 32) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar + 1)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::);
        const uint32_t bytes = (uint32_t)H * (uint32_t)H * 2u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(s_bar + 1)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sB)), "l"(f.w1p + (long long)z * f.w1p_zs), "r"(bytes), "r"(smem_u32(s_bar + 1)) : "memory");
    }
    // layer 2 mapping as in the forward kernel: lane = (row of an 8-row block, 8-column group of a 32-column step)
    const int rsub = lane & 7, grp = lane >> 3;
    float d0[FUSED_RB], d1[FUSED_RB];
    uint4 mlo[FUSED_RB], mhi[FUSED_RB];
#pragma unroll
    for (int rb = 0; rb < FUSED_RB; ++rb) {
        const int m = m0 + warp * (8 * FUSED_RB) + rb * 8 + rsub;
        const bool lv = m < f.B;
        const float* dout = f.DOut + (long long)z * f.B * f.nout + (long long)m * f.nout;
        d1[rb] = (lv && f.nout == 2) ? dout[1] : 0.f;
        if (f.seed == 0 || !lv) {
            d0[rb] = lv ? dout[0] : 0.f;
        } else if (f.seed == 1) {
            const float invB = 1.0f / (float)f.B;
            const float qt = fminf(f.oqt[m], f.oqt[f.B + m]) - f.alpha[0] * f.logp_next[m];
            const float y = fmaf(f.msk[m] * f.gamma, qt, f.rew[m]);
            d0[rb] = 2.0f * (f.oq[(long long)z * f.B + m] - y) * invB;
        } else {
            const float invB = 1.0f / (float)f.B;
            const float qa = f.oq[(long long)z * f.B + m], qb = f.oq[(long long)(1 - z) * f.B + m];
            d0[rb] = qa < qb ? -invB : (qa == qb ? -0.5f * invB : 0.f);
        }
        const uint4* mp = reinterpret_cast<const uint4*>(f.M2 + ((long long)z * f.B + m) * 8);
        mlo[rb] = lv ? mp[0] : make_uint4(0u, 0u, 0u, 0u);
        mhi[rb] = lv ? mp[1] : make_uint4(0u, 0u, 0u, 0u);
    }
    // epilogue mask (h1 > 0) of this thread's accumulator row
    const int row = tid & (TCG_M - 1), half = tid >> 7, kh = H / FUSED_NQ;      // (`half`: this thread's column part)
    const int m = m0 + row;
    const bool live = m < f.B;
    uint4 m1lo = make_uint4(0u, 0u, 0u, 0u), m1hi = m1lo;
    if (live) {
        const uint4* mp = reinterpret_cast<const uint4*>(f.M1 + ((long long)z * f.B + m) * 8);
        m1lo = mp[0]; m1hi = mp[1];
    }
    {   // small operands: all loads of a thread in flight together (see the forward kernel)
        constexpr int Q2 = (2 * 256 + FUSED_NT - 1) / FUSED_NT;
        float t2[Q2], tx = 0.f;
        const int n2 = H * f.nout;                                       // <= 512
#pragma unroll
        for (int q = 0; q < Q2; ++q) { const int i = tid + q * FUSED_NT; t2[q] = i < n2 ? W[f.l2 + i] : 0.f; }
        if (tid < H && f.DX) tx = W[f.l0 + (long long)f.dx_row * H + tid];
#pragma unroll
        for (int q = 0; q < Q2; ++q) {
            const int i = tid + q * FUSED_NT;
            if (i < n2) {
                const int k = i / f.nout, j = i % f.nout;
                s_w2[2 * k + j] = t2[q];
                if (f.nout == 1) s_w2[2 * k + 1] = 0.f;
            }
        }
        if (tid < H) s_wx[tid] = tx;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    const uint32_t tmem_base = *s_tmem;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(H >> 3) << 17) | ((uint32_t)(TCG_M >> 4) << 24);

    // layer 2: d h2 = (d out . W2^T) where h2 > 0
#pragma unroll
    for (int rb = 0; rb < FUSED_RB; ++rb) {
        const int r2 = warp * (8 * FUSED_RB) + rb * 8 + rsub, m2 = m0 + r2;
        float* dh2 = f.DH2 ? f.DH2 + ((long long)z * f.B + m2) * H : nullptr;
        for (int it = 0; it < (H >> 5); ++it) {
            const int k0 = it * 32 + grp * 8;
            const uint32_t bits = mask_word(mlo[rb], mhi[rb], it) >> (8 * grp);
            float a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float2 w = reinterpret_cast<const float2*>(s_w2)[k0 + j];
                float acc = fmaf(d0[rb], w.x, 0.f);
                if (f.nout == 2) acc = fmaf(d1[rb], w.y, acc);
                a[j] = (bits >> j & 1u) ? acc : 0.f;
            }
            if (m2 < f.B && dh2) {
                *reinterpret_cast<float4*>(dh2 + k0) = make_float4(a[0], a[1], a[2], a[3]);
                *reinterpret_cast<float4*>(dh2 + k0 + 4) = make_float4(a[4], a[5], a[6], a[7]);
            }
            uint4 v;
            v.x = pack_bf16x2(a[0], a[1]); v.y = pack_bf16x2(a[2], a[3]); v.z = pack_bf16x2(a[4], a[5]); v.w = pack_bf16x2(a[6], a[7]);
            unsigned char* core = sA + (size_t)(r2 >> 3) * sbo + (size_t)(k0 >> 3) * lbo + (r2 & 7) * 16;
            *reinterpret_cast<uint4*>(core) = v;
            if (f.IMG2) {
                image_store_warp(f.IMG2 + (long long)z * f.img_zs, (int)gridDim.x * 2, it * 32, m2 - rsub, smem_u32(core), lane);
                // column sums of this warp's 8 rows (fp32, fixed tree), for the bias row of d W1
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float c = a[j];
                    c += __shfl_xor_sync(0xffffffffu, c, 1); c += __shfl_xor_sync(0xffffffffu, c, 2); c += __shfl_xor_sync(0xffffffffu, c, 4);
                    if (rsub == 0) s_cs[(warp * FUSED_RB + rb) * H + k0 + j] = c;
                }
            }
        }
    }
    if (f.IMG2) {                                                       // 8-row sums -> the CTA's 128-row sums, ascending row order
        __syncthreads();
        for (int n = tid; n < H; n += FUSED_NT) {
            float c = 0.f;
            for (int q = 0; q < FUSED_NW * FUSED_RB; ++q) c += s_cs[q * H + n];
            f.colsum[((long long)z * gridDim.x + blockIdx.x) * H + n] = c;
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    // layer 1 on the tensor cores: D[128 x H(k)] = d h2 [128 x H(n)] . W1^T [H(n) x H(k)]
    if (tid == 0) {
        mbar_wait(smem_u32(s_bar + 1), 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        const uint64_t da = umma_desc_kmajor_noswizzle(smem_u32(sA), lbo, sbo), db = umma_desc_kmajor_noswizzle(smem_u32(sB), lbo, sbo);
        for (int ks = 0; ks < (H >> 4); ++ks) {
            const uint64_t adv = (uint64_t)((ks * 2 * lbo) >> 4);
            const uint32_t acc = ks > 0 ? 1u : 0u;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tmem_base), "l"(da + adv), "l"(db + adv), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(s_bar)) : "memory");
    }
    mbar_wait(smem_u32(s_bar), 0u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    // epilogue: d h1 = D where h1 > 0 -> HBM through the warp's staging tile; d x accumulated on the way
    float dx = 0.f;
    float* stg = reinterpret_cast<float*>(sA) + warp * 32 * FUSED_STG_STRIDE;
    const int wrow0 = m0 + (warp & 3) * 32;
    float* dh1w = f.DH1 ? f.DH1 + ((long long)z * f.B + wrow0) * H : nullptr;
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (int c0 = half * kh; c0 < (half + 1) * kh; c0 += 16) {
        uint32_t r[16];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                       "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(taddr + (uint32_t)c0));
        const uint32_t bits = mask_word(m1lo, m1hi, c0 >> 5) >> (c0 & 31);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        float g[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            g[j] = (bits >> j & 1u) ? __uint_as_float(r[j]) : 0.f;
            dx = fmaf(g[j], s_wx[c0 + j], dx);
        }
        if (dh1w || f.IMG1) staged_store16(stg, lane, g, dh1w ? dh1w + c0 : nullptr, H, f.B - wrow0,
                                 f.IMG1 ? f.IMG1 + (long long)z * f.img_zs + ((size_t)(c0 >> 7) * (gridDim.x * 2) + (wrow0 >> 6)) * 16384 : nullptr, c0, wrow0 & 63);
    }
    if (half > 0) s_part[(half - 1) * TCG_M + row] = dx;
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    if (half == 0 && live && f.DX) {
#pragma unroll
        for (int q = 0; q < FUSED_NQ - 1; ++q) dx += s_part[q * TCG_M + row];       // fixed order
        f.DX[(long long)z * f.B + m] = dx;
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(f.tmem_cols));
}

// ---- layer 1's weight gradient from the operand images ----------------------------------------------------------------------
// d W1[k][n] = sum over the batch of h1[b][k] d h2[b][n]: both operands are bf16 tile images with the batch as K (written by
// the fused kernels, above), so a CTA's whole k range is fetched with two bulk copies per 64-row chunk -- all of them in
// flight at once -- instead of being converted through registers chunk after chunk.  Grid (n tiles, m tiles, Z x splits),
// 128 x 128 output tile, fp32 partial tiles [z][split][H][H] summed in split order by dw1_reduce_kernel.
constexpr int DWI_STAGES = 4;
struct DwImg {
    const unsigned char *A, *Bm; long long a_zs, b_zs;   // images of h1 and d h2, bytes between networks
    int n_chunks, splits, H, M;                           // 64-row chunks per network; M = rows of the product (features of A)
    float* partial;                                       // [Z][splits][M][H]
};
__global__ void __launch_bounds__(256, 1) dw_img_tc_kernel(const DwImg t) {
    extern __shared__ __align__(128) unsigned char dwi_smem[];   // DWI_STAGES x (A 16 KB | B 16 KB); after the MMAs: staging tiles
    __shared__ uint64_t s_full[DWI_STAGES], s_done;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int z = blockIdx.z / t.splits, sp = blockIdx.z % t.splits;
    const int per = (t.n_chunks + t.splits - 1) / t.splits;
    const int c_lo = sp * per, c_hi = min(t.n_chunks, c_lo + per), n_my = max(0, c_hi - c_lo);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(128));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 32) {
        for (int i = 0; i < DWI_STAGES; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_full[i])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_done)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::);
        // a CTA's k range is at most DWI_STAGES chunks (the host picks the splits so): every block is requested at once
        const unsigned char* a = t.A + (long long)z * t.a_zs + ((size_t)blockIdx.y * t.n_chunks + c_lo) * 16384;
        const unsigned char* b = t.Bm + (long long)z * t.b_zs + ((size_t)blockIdx.x * t.n_chunks + c_lo) * 16384;
        for (int i = 0; i < n_my; ++i) {
            const uint32_t bar = smem_u32(&s_full[i]), dstA = smem_u32(dwi_smem + (size_t)i * 32768), dstB = dstA + 16384;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(32768u) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dstA), "l"(a + (size_t)i * 16384), "r"(16384u), "r"(bar) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dstB), "l"(b + (size_t)i * 16384), "r"(16384u), "r"(bar) : "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    const uint32_t tmem_base = s_tmem;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(TCG_M >> 4) << 24);
    if (tid == 0) {
        for (int i = 0; i < n_my; ++i) {
            mbar_wait(smem_u32(&s_full[i]), 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::);
            const uint32_t sa = smem_u32(dwi_smem + (size_t)i * 32768);
            const uint64_t da = umma_desc_kmajor_noswizzle(sa, 128u, 1024u), db = umma_desc_kmajor_noswizzle(sa + 16384u, 128u, 1024u);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint64_t adv = (uint64_t)((ks * 256) >> 4);
                const uint32_t acc = (i > 0 || ks > 0) ? 1u : 0u;
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             ::"r"(tmem_base), "l"(da + adv), "l"(db + adv), "r"(idesc), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_done)) : "memory");
    }
    mbar_wait(smem_u32(&s_done), 0u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    // accumulator rows (thread = TMEM lane, the two thread halves take the two column halves) -> the partial tile, through the
    // warp's staging tile (8 rows x 64 contiguous bytes per store)
    const int half = tid >> 7;
    float* stg = reinterpret_cast<float*>(dwi_smem) + warp * 32 * FUSED_STG_STRIDE;
    const int H = t.H, m0 = blockIdx.y * TCG_M + (warp & 3) * 32, n0 = blockIdx.x * 128;
    float* P = t.partial + (((long long)z * t.splits + sp) * t.M + m0) * H + n0;
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (int c0 = half * 64; c0 < (half + 1) * 64; c0 += 16) {
        uint32_t r[16];
        float v[16];
        if (n_my > 0) {
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                           "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                         : "r"(taddr + (uint32_t)c0));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = n_my > 0 ? __uint_as_float(r[j]) : 0.f;
        if (n0 + c0 < H) staged_store16(stg, lane, v, P + c0, H, t.M - m0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128));
}
// grad[z][m][n] (augmented layer [M (+ 1)][H]) = sum over the splits, in split order, of the partial tiles (m < M); with
// `colsum`: one more row, the bias row of layer 1 = sum over the row tiles, in ascending order, of the fused backward
// kernel's column sums
__global__ void __launch_bounds__(256) dw1_reduce_kernel(const DwImg t, float* grad, const long long g_zs, const float* colsum, const int row_tiles) {
    const int H = t.H, z = blockIdx.y;
    const long long mh = (long long)t.M * H;
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= mh + (colsum ? H : 0)) return;
    float acc = 0.f;
    if (i < mh) {
        const float* pp = t.partial + (long long)z * t.splits * mh + i;
        for (int q0 = 0; q0 < t.splits; q0 += 8) {
            float x[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) x[q] = q0 + q < t.splits ? pp[(long long)(q0 + q) * mh] : 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) acc += x[q];
        }
    } else {
        const int n = (int)(i - mh);
        for (int q = 0; q < row_tiles; ++q) acc += colsum[((long long)z * row_tiles + q) * H + n];
    }
    grad[(long long)z * g_zs + i] = acc;
}

// ---- workspace ------------------------------------------------------------------------------------------------------
struct Net { long long l0, l1, l2; };   // float offsets of the three augmented layer blocks inside a network block
struct Layout {
    int H, B;
    long long P_pol, P_q;               // floats per policy / critic block
    Net pol, q;
    // float offsets inside the workspace
    long long o_pol, o_q, o_qt, o_log_alpha, o_alpha, o_grad, o_m, o_v, n_train;   // n_train = P_pol + 2 P_q + 1 (log_alpha)
    long long o_xs, o_xs2, o_xsa, o_xs2a, o_xspi, o_r, o_mask;
    long long o_hp1, o_hp2, o_op, o_hq1, o_hq2, o_oq, o_oqt, o_hqt1, o_hqt2;   // hqt*: scratch activations of the target critics
    long long o_doq, o_dhq2, o_dhq1, o_da, o_dop, o_dhp2, o_dhp1;
    long long o_eps_next, o_eps_cur, o_logp_next, o_logp_cur, o_idx, o_losses, o_updates, o_obs_scale, o_part, o_w1p, o_mhp, o_mhq, o_img_hp1, o_img_hq1, o_img_dhp2, o_img_dhq2, o_img_xp, o_img_xq, o_img_dhp1, o_img_dhq1, o_colsum, total;
    int tc;                             // large-batch products on tcgen05 (gemm_tc_kernel)
    int dw_img;                         // fused path: layer 1's weight gradient from bf16 tile images (dw_img_tc_kernel); SIT_SAC_DW_IMG=0: gemm_tc_kernel
    int fused_fwd;                      // tc: a network's forward as ONE kernel (mlp_fwd_tc_kernel); SIT_SAC_FUSED_FWD=0 keeps the three products
    int ldh;                            // leading dimension of the hidden activations [B][ldh]: H weights' inputs + the ones
                                        // column (H + 1), padded to a multiple of 4 floats on the tensor-core path so that
                                        // every row starts 16-byte aligned (128-bit operand loads)
};

static void make_layout(Layout& L, int H, int B, int tc) {
    L.H = H; L.B = B; L.tc = tc; L.ldh = tc ? H + 4 : H + 1;
    L.fused_fwd = tc && H % (32 * FUSED_NQ) == 0;   // column quarters of whole 32-bit mask words: H = 128 or 256
    if (const char* e = std::getenv("SIT_SAC_FUSED_FWD")) if (std::atoi(e) == 0) L.fused_fwd = 0;
    L.pol.l0 = 0; L.pol.l1 = (long long)(S + 1) * H; L.pol.l2 = L.pol.l1 + (long long)(H + 1) * H;
    L.P_pol = L.pol.l2 + (long long)(H + 1) * 2;
    L.q.l0 = 0; L.q.l1 = (long long)(SA + 1) * H; L.q.l2 = L.q.l1 + (long long)(H + 1) * H;
    L.P_q = L.q.l2 + (long long)(H + 1) * 1;
    long long o = 0;
    auto take = [&](long long n) { const long long at = o; o += (n + 63) / 64 * 64; return at; };   // 256-byte aligned
    // trainable block is contiguous: policy | q1 | q2 | log_alpha  (grad / Adam moments mirror it)
    L.o_pol = o; o += L.P_pol; L.o_q = o; o += 2 * L.P_q; L.o_log_alpha = o; o += 1;
    L.n_train = L.P_pol + 2 * L.P_q + 1;
    o = (o + 63) / 64 * 64;
    L.o_qt = take(2 * L.P_q); L.o_alpha = take(1);
    L.o_grad = take(L.n_train); L.o_m = take(L.n_train); L.o_v = take(L.n_train);
    const long long b = B;
    L.o_xs = take(b * (S + 1)); L.o_xs2 = take(b * (S + 1));
    L.o_xsa = take(b * (SA + 1)); L.o_xs2a = take(b * (SA + 1)); L.o_xspi = take(b * (SA + 1));
    L.o_r = take(b); L.o_mask = take(b);
    L.o_hp1 = take(b * L.ldh); L.o_hp2 = take(b * L.ldh); L.o_op = take(b * 2);
    L.o_hq1 = take(2 * b * L.ldh); L.o_hq2 = take(2 * b * L.ldh); L.o_oq = take(2 * b); L.o_oqt = take(2 * b);
    L.o_hqt1 = take(2 * b * L.ldh); L.o_hqt2 = take(2 * b * L.ldh);   // own buffers: the target pass runs beside / after the critics' forward on (s, a), whose activations the backward needs
    L.o_doq = take(2 * b); L.o_dhq2 = take(2 * b * H); L.o_dhq1 = take(2 * b * H); L.o_da = take(2 * b);
    L.o_dop = take(b * 2); L.o_dhp2 = take(b * H); L.o_dhp1 = take(b * H);
    L.o_eps_next = take(b); L.o_eps_cur = take(b); L.o_logp_next = take(b); L.o_logp_cur = take(b);
    L.o_idx = take(b); L.o_losses = take(SIT_SAC_N_LOSS); L.o_updates = take(2);   // updates: one unsigned long long
    L.o_obs_scale = take(S);
    L.o_part = take(tc ? 3 * 2LL * TCG_MAX_SPLITS * (H + 1) * H : 0);   // split-k partial tiles [z][split][M][N], one region per concurrent weight-gradient product
    L.o_w1p = take(tc ? (long long)SAC_W1_IMAGES * H * H / 2 : 0);   // bf16 K-major images of W1 (pack_w1_nets_kernel)
    L.o_mhp = take(tc ? 2LL * b * 8 : 0);        // sign bits of the policy's h1 | h2 (fused forward -> fused backward), 8 words per row
    L.o_mhq = take(tc ? 2LL * 2 * b * 8 : 0);    // ... of the critics' h1 [2][B] | h2 [2][B]
    // operand images of layer 1's weight gradient d W1 = h1^T . d h2 (contraction over the batch): bf16, [z][128-feature
    // tile][64-row chunk] blocks of 16 KB in the UMMA canonical K-major layout (K = batch), written by the fused forward
    // (h1) and backward (d h2) kernels, fetched by dw_img_tc_kernel with one bulk copy per block
    // from batch 256 (below, the image stores and the two reduce launches cost more than the products save: batch 64 10 600 ->
    // 9 550 updates/s) up to the batch whose k range fits the stages of a CTA at the largest split count (8192)
    L.dw_img = L.fused_fwd && B >= 256 && (B + TCG_M - 1) / TCG_M * 2 <= DWI_STAGES * TCG_MAX_SPLITS;
    if (const char* e = std::getenv("SIT_SAC_DW_IMG")) if (std::atoi(e) == 0) L.dw_img = 0;
    {
        const long long rows = (b + TCG_M - 1) / TCG_M * TCG_M, img = L.dw_img ? rows * H * 2 / 4 : 0;   // floats per network
        L.o_img_hp1 = take(img); L.o_img_hq1 = take(2 * img); L.o_img_dhp2 = take(img); L.o_img_dhq2 = take(2 * img);
        // layer 0's weight gradient d W0 = x^T . d h1: the inputs as one 128-feature tile (11 / 12 of them used, the rest stays
        // zero from sit_sac_init), d h1 like d h2
        const long long ximg = L.dw_img ? rows * 128 * 2 / 4 : 0;
        L.o_img_xp = take(ximg); L.o_img_xq = take(ximg); L.o_img_dhp1 = take(img); L.o_img_dhq1 = take(2 * img);
        L.o_colsum = take(L.dw_img ? 2 * (rows / TCG_M) * H : 0);   // per-CTA column sums of d h2 (the bias row of d W1)
    }
    L.total = o;
}

struct Hyper { float gamma, tau, lr, alpha0; int auto_alpha, target_interval; uint64_t seed; };

// ---- element-wise kernels -------------------------------------------------------------------------------------------
// Glorot-uniform init of one augmented layer [K+1][N] (weights U(-l, l), l = sqrt(6 / (K + N)); bias row 0): mlp.py:13-51
__global__ void init_layer_kernel(float* w, int K, int N, uint64_t seed, uint32_t stream_id) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)(K + 1) * N) return;
    float v = 0.f;
    if (i < (long long)K * N) {
        uint32_t x[4];
        philox4x32_10((uint32_t)i, (uint32_t)(i >> 32), stream_id, 0x5ac1u, (uint32_t)seed, (uint32_t)(seed >> 32), x);
        const float u = ((float)x[0] + 0.5f) * (1.0f / 4294967296.0f);
        v = (2.f * u - 1.f) * sqrtf(6.0f / (float)(K + N));
    }
    w[i] = v;
}

struct ObsScale { float v[SIT_N_OBS]; };
__global__ void obs_scale_kernel(float* p, const ObsScale sc) {
    if (threadIdx.x < SIT_N_OBS) p[threadIdx.x] = sc.v[threadIdx.x] != 0.0f ? sc.v[threadIdx.x] : 1.0f;
}
__global__ void fill_kernel(float* p, long long n, float v) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
// ones column of an activation buffer [rows][ld] (written once: no kernel ever touches that column again)
__global__ void ones_col_kernel(float* p, long long rows, int ld, int col) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < rows) p[i * ld + col] = 1.0f;
}

// draw the batch: uniform row indices over the valid part of the replay ring (memory.sample, main_ast.py:354), gather
// (s, a, r, s', mask) into the activation buffers, draw the two standard normals of this update for every row
__global__ void batch_kernel(float* ws, const Layout L, const Hyper hp, const float* replay, long long capacity,
                             const unsigned long long* n_rows) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0) {
        float* ls = ws + L.o_losses;
        for (int i = 0; i < SIT_SAC_N_LOSS; ++i) ls[i] = 0.f;
    }
    if (b >= L.B) return;
    const unsigned long long upd = *reinterpret_cast<const unsigned long long*>(ws + L.o_updates);
    unsigned long long nv = n_rows ? *n_rows : (unsigned long long)capacity;
    if (nv > (unsigned long long)capacity) nv = (unsigned long long)capacity;
    if (nv == 0) nv = 1;
    uint32_t x[4];
    philox4x32_10((uint32_t)b, 0x5ac0u, (uint32_t)upd, (uint32_t)(upd >> 32), (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), x);
    const long long row = (long long)((((unsigned long long)x[0] << 32) | x[1]) % nv);
    reinterpret_cast<int32_t*>(ws + L.o_idx)[b] = (int32_t)row;
    const float u1 = ((float)x[2] + 0.5f) * (1.0f / 4294967296.0f), u2 = ((float)x[3] + 0.5f) * (1.0f / 4294967296.0f);
    const float rad = sqrtf(-2.0f * logf(u1));
    float sn, cs;
    sincospif(2.0f * u2, &sn, &cs);
    ws[L.o_eps_next + b] = rad * cs;
    ws[L.o_eps_cur + b] = rad * sn;
    const float* tr = replay + row * SIT_TR_FLOATS;
    float* xs = ws + L.o_xs + (long long)b * (S + 1);
    float* xs2 = ws + L.o_xs2 + (long long)b * (S + 1);
    float* xsa = ws + L.o_xsa + (long long)b * (SA + 1);
    float* xs2a = ws + L.o_xs2a + (long long)b * (SA + 1);
    float* xspi = ws + L.o_xspi + (long long)b * (SA + 1);
#pragma unroll
    for (int c = 0; c < S; ++c) {
        const float sc = ws[L.o_obs_scale + c];            // SitSacConfig.obs_scale: the networks see scaled observations
        const float s = tr[c] * sc, s2 = tr[12 + c] * sc;
        xs[c] = s; xsa[c] = s; xspi[c] = s; xs2[c] = s2; xs2a[c] = s2;
    }
    xsa[S] = tr[10];
    ws[L.o_r + b] = tr[11];
    ws[L.o_mask + b] = tr[22];
}

// tanh-Gaussian head (gaussian_policy.py:116-117,143-148; normal.py:86-112): a = tanh(mu + sigma eps) written into the
// action column of a critic input buffer, log pi = log N(x; mu, sigma) - log(1 - a^2 + 1e-6)
__global__ void head_kernel(float* ws, const Layout L, long long o_eps, long long o_xa, long long o_logp) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= L.B) return;
    const float mu = ws[L.o_op + 2 * b], ls = fminf(fmaxf(ws[L.o_op + 2 * b + 1], LOG_SIG_MIN), LOG_SIG_MAX);
    const float eps = ws[o_eps + b];
    const float a = tanhf(fmaf(expf(ls), eps, mu));
    ws[o_xa + (long long)b * (SA + 1) + S] = a;
    ws[o_logp + b] = -0.5f * eps * eps - ls - 0.91893853320467274f - logf(1.0f - a * a + SQUASH_EPS);
}

// deterministic block reduction helper (1024 threads, fixed tree)
__device__ __forceinline__ float block_sum_1024(float v, float* sm) {
    const int t = threadIdx.x;
    sm[t] = v;
    __syncthreads();
    for (int s = 512; s > 0; s >>= 1) {
        if (t < s) sm[t] += sm[t + s];
        __syncthreads();
    }
    const float r = sm[0];
    __syncthreads();
    return r;
}

// y = r + mask gamma (min(Q1t, Q2t)(s', a') - alpha log pi');  d mse / d q_z = 2 (q_z - y) / B.  One CTA.
__global__ void __launch_bounds__(1024) critic_loss_kernel(float* ws, const Layout L, const Hyper hp) {
    __shared__ float sm[1024];
    const float alpha = ws[L.o_alpha];
    const float invB = 1.0f / (float)L.B;
    float l1 = 0.f, l2 = 0.f, sq = 0.f;
    for (int b = threadIdx.x; b < L.B; b += 1024) {
        const float qt = fminf(ws[L.o_oqt + b], ws[L.o_oqt + L.B + b]) - alpha * ws[L.o_logp_next + b];
        const float y = fmaf(ws[L.o_mask + b] * hp.gamma, qt, ws[L.o_r + b]);
        const float d1 = ws[L.o_oq + b] - y, d2 = ws[L.o_oq + L.B + b] - y;
        ws[L.o_doq + b] = 2.0f * d1 * invB;
        ws[L.o_doq + L.B + b] = 2.0f * d2 * invB;
        l1 += d1 * d1; l2 += d2 * d2; sq += ws[L.o_oq + b];
    }
    l1 = block_sum_1024(l1, sm); l2 = block_sum_1024(l2, sm); sq = block_sum_1024(sq, sm);
    if (threadIdx.x == 0) {
        ws[L.o_losses + SIT_SAC_LOSS_Q1] = l1 * invB; ws[L.o_losses + SIT_SAC_LOSS_Q2] = l2 * invB;
        ws[L.o_losses + SIT_SAC_MEAN_Q] = sq * invB;
    }
}

// policy loss = mean(alpha log pi - min(Q1, Q2)(s, a~)): seeds d loss / d q_z (ties split evenly like torch.min) and
// accumulates the loss and mean log pi.  One CTA.
__global__ void __launch_bounds__(1024) policy_seed_kernel(float* ws, const Layout L) {
    __shared__ float sm[1024];
    const float alpha = ws[L.o_alpha];
    const float invB = 1.0f / (float)L.B;
    float pl = 0.f, lp = 0.f;
    for (int b = threadIdx.x; b < L.B; b += 1024) {
        const float q1 = ws[L.o_oq + b], q2 = ws[L.o_oq + L.B + b], logp = ws[L.o_logp_cur + b];
        ws[L.o_doq + b] = q1 < q2 ? -invB : (q1 == q2 ? -0.5f * invB : 0.f);
        ws[L.o_doq + L.B + b] = q2 < q1 ? -invB : (q1 == q2 ? -0.5f * invB : 0.f);
        pl += alpha * logp - fminf(q1, q2);
        lp += logp;
    }
    pl = block_sum_1024(pl, sm); lp = block_sum_1024(lp, sm);
    if (threadIdx.x == 0) { ws[L.o_losses + SIT_SAC_LOSS_POLICY] = pl * invB; ws[L.o_losses + SIT_SAC_MEAN_LOGPI] = lp * invB; }
}

// d loss / d (mu, log sigma) from d loss / d a (through both critics) and the entropy term
__global__ void policy_grad_kernel(float* ws, const Layout L) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= L.B) return;
    const float alpha = ws[L.o_alpha], invB = 1.0f / (float)L.B;
    const float ls_raw = ws[L.o_op + 2 * b + 1];
    const float ls = fminf(fmaxf(ls_raw, LOG_SIG_MIN), LOG_SIG_MAX);
    const float eps = ws[L.o_eps_cur + b];
    const float a = ws[L.o_xspi + (long long)b * (SA + 1) + S];
    const float one_m = 1.0f - a * a;
    const float dq_da = ws[L.o_da + b] + ws[L.o_da + L.B + b];
    // log pi = -eps^2/2 - ls - c - log(1 - a^2 + e),  a = tanh(x),  x = mu + exp(ls) eps
    const float gx = alpha * invB * (2.0f * a * one_m / (one_m + SQUASH_EPS)) + dq_da * one_m;
    const bool in_range = ls_raw >= LOG_SIG_MIN && ls_raw <= LOG_SIG_MAX;   // clip passes no gradient outside
    ws[L.o_dop + 2 * b] = gx;
    ws[L.o_dop + 2 * b + 1] = in_range ? fmaf(gx, expf(ls) * eps, -alpha * invB) : 0.f;
}

// Adam step (torch.optim.Adam defaults, main_ast.py:51 lr) on a parameter range, optional fused soft target update
// target <- (1 - tau) target + tau param (main_ast.py:43-44,84)
// pack_images != 0 (the critics' step on the fused tensor-core path): the new W1 entries also go, as bf16, into the forward
// and by-k images of q1 / q2 (images 1, 2 and 6, 7 of pack_w1_nets_kernel: same rounding, same positions), so the forward
// that follows needs no pack launch
__global__ void adam_kernel(float* ws, const Layout L, const Hyper hp, long long first, long long count, long long target_delta,
                            int soft, int pack_images = 0) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const unsigned long long upd = *reinterpret_cast<const unsigned long long*>(ws + L.o_updates);
    const float t = (float)(upd + 1);
    const float bc1 = 1.0f - powf(ADAM_B1, t), bc2 = 1.0f - powf(ADAM_B2, t);
    const long long j = first + i;                       // index inside the trainable block
    const float g = ws[L.o_grad + j];
    const float m = ADAM_B1 * ws[L.o_m + j] + (1.0f - ADAM_B1) * g;
    const float v = ADAM_B2 * ws[L.o_v + j] + (1.0f - ADAM_B2) * g * g;
    ws[L.o_m + j] = m; ws[L.o_v + j] = v;
    const float p = ws[L.o_pol + j] - (hp.lr / bc1) * m / (sqrtf(v) / sqrtf(bc2) + ADAM_EPS);
    ws[L.o_pol + j] = p;
    if (pack_images) {
        const int H = L.H;
        const long long net = i / L.P_q, e = i % L.P_q - L.q.l1;
        if (e >= 0 && e < (long long)H * H) {
            const int k = (int)(e / H), n = (int)(e % H);
            unsigned short hb;
            asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(hb) : "f"(p));
            unsigned char* img = reinterpret_cast<unsigned char*>(ws + L.o_w1p);
            const size_t sbo = 16u * (size_t)H, lbo = 128u, isz = (size_t)H * H * 2;
            *reinterpret_cast<unsigned short*>(img + (1 + net) * isz + (size_t)(n >> 3) * sbo + (size_t)(k >> 3) * lbo + (n & 7) * 16 + (k & 7) * 2) = hb;
            *reinterpret_cast<unsigned short*>(img + (6 + net) * isz + (size_t)(k >> 3) * sbo + (size_t)(n >> 3) * lbo + (k & 7) * 16 + (n & 7) * 2) = hb;
        }
    }
    if (soft && (upd % (unsigned long long)hp.target_interval) == 0) {
        const long long k = L.o_pol + j + target_delta;
        ws[k] = fmaf(hp.tau, p - ws[k], ws[k]);
    }
}

// entropy temperature: alpha loss = -mean(log_alpha (log pi + target_entropy)); Adam on log_alpha; alpha = exp(log_alpha)
// from the next update on; closes the update (counter + 1).  One thread.
__global__ void alpha_kernel(float* ws, const Layout L, const Hyper hp, float* losses_out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long* updp = reinterpret_cast<unsigned long long*>(ws + L.o_updates);
    const unsigned long long upd = *updp;
    float* ls = ws + L.o_losses;
    const float mean_logp = ls[SIT_SAC_MEAN_LOGPI];
    if (hp.auto_alpha) {
        const float log_alpha = ws[L.o_log_alpha];
        ls[SIT_SAC_LOSS_ALPHA] = -log_alpha * (mean_logp + TARGET_ENTROPY);
        const float g = -(mean_logp + TARGET_ENTROPY);
        const long long j = L.n_train - 1;
        const float t = (float)(upd + 1);
        const float m = ADAM_B1 * ws[L.o_m + j] + (1.0f - ADAM_B1) * g;
        const float v = ADAM_B2 * ws[L.o_v + j] + (1.0f - ADAM_B2) * g * g;
        ws[L.o_m + j] = m; ws[L.o_v + j] = v;
        const float bc1 = 1.0f - powf(ADAM_B1, t), bc2 = 1.0f - powf(ADAM_B2, t);
        const float nl = log_alpha - (hp.lr / bc1) * m / (sqrtf(v) / sqrtf(bc2) + ADAM_EPS);
        ws[L.o_log_alpha] = nl;
        ws[L.o_alpha] = expf(nl);
    }
    ls[SIT_SAC_ALPHA] = ws[L.o_alpha];
    if (losses_out)
        for (int i = 0; i < SIT_SAC_N_LOSS; ++i) losses_out[i] = ls[i];
    *updp = upd + 1;
}

}  // namespace sac

// =================================================================================================
// Host side of the learner
// =================================================================================================
struct SitSac_ {
    SitSacConfig cfg;
    int device;
    sac::Layout L;
    sac::Hyper hp;
    cudaGraphExec_t graph; std::string graph_key;
    // second branch of the update on the tensor-core path: the three weight-gradient products of a backward pass run beside
    // the input-gradient chain (they only depend on it from the side), captured into the same CUDA graph
    cudaStream_t side; cudaEvent_t ev[4];
    // third branch: the critics' forward on (s, a) beside the target computation, and the policy's forward on s beside the
    // critics' backward + Adam (independent until the joins below)
    cudaStream_t side2; cudaEvent_t ev2[4];
    // fourth: layer 0's weight-gradient product beside layer 1's (fused backward)
    cudaStream_t side3; cudaEvent_t ev3[4];     // [2], [3]: the W1 images of an update packed beside its batch kernel
};

namespace sac {

static size_t bwd_smem_bytes(int H) { return (size_t)TCG_M * H * 2 + (size_t)H * H * 2 + (size_t)12 * H + 512 * (FUSED_NQ - 1) + 32 + (size_t)FUSED_NW * FUSED_RB * H * 4; }
static size_t fwd_smem_bytes(int H, int kin1) { return (size_t)TCG_M * H * 2 + (size_t)H * H * 2 + (size_t)12 * H + 16 + 1024 * (FUSED_NQ - 1) + 32 + (size_t)kin1 * H * 4; }

struct Launcher {
    cudaStream_t st;
    float* ws;
    const Layout& L;
    cudaStream_t side;           // nullptr: everything on `st`
    cudaEvent_t* ev;
    mutable int branch;          // 0: launch on st; 1 .. 3: launch on `side` (3: on `side_b` when there is one), split-k partials in region branch - 1
    const Hyper* hp = nullptr;       // for the loss kernels the fused backward moves to the side stream
    cudaStream_t side_b = nullptr;   // second side stream of the fused backward: layer 0's weight gradient beside layer 1's
    cudaEvent_t* ev_b = nullptr;
    // C[z] (M x N, ldc) = epi(A[z] . B[z]);  A(m,k) = A[m a_rs + k a_cs], B(k,n) = B[k b_rs + n b_cs]
    void gemm(long long oA, int a_rs, int a_cs, long long a_zs, long long oB, int b_rs, int b_cs, long long b_zs, long long oC,
              int ldc, long long c_zs, int M, int N, int K, int Z, int epi, long long oH = 0, int ldh = 0, long long h_zs = 0) const {
        Gemm g;
        g.A = ws + oA; g.B = ws + oB; g.C = ws + oC; g.Hm = ws + oH;
        g.M = M; g.N = N; g.K = K; g.a_rs = a_rs; g.a_cs = a_cs; g.b_rs = b_rs; g.b_cs = b_cs; g.ldc = ldc; g.ldh = ldh;
        g.a_zs = a_zs; g.b_zs = b_zs; g.c_zs = c_zs; g.h_zs = h_zs; g.epi = epi;
        cudaStream_t st = branch > 0 ? ((branch == 3 && side_b) ? side_b : side) : this->st;      // the branch this product belongs to
        if (L.tc && K <= 16) {
            gemm_thin_k_kernel<<<dim3((M + THIN_ROWS - 1) / THIN_ROWS, 1, Z), 256, 0, st>>>(g);
            return;
        }
        if (L.tc && N <= 4 && a_cs == 1 && K >= 128) {
            gemm_skinny_n_kernel<<<dim3((M + 7) / 8, 1, Z), 256, 0, st>>>(g);
            return;
        }
        if (L.tc && K >= 128) {      // hidden-layer products and every contraction over the batch: tensor cores
            GemmTC t;
            t.g = g;
            t.bn = N > 64 ? 128 : (N > 32 ? 64 : (N > 16 ? 32 : 16));
            t.tmem_cols = t.bn < 32 ? 32 : t.bn;
            const int n_chunks = (K + TCG_KC - 1) / TCG_KC;
            int splits = K >= 1024 ? std::min(TCG_MAX_SPLITS, std::max(1, n_chunks / 4)) : 1;
            {   // no more CTAs than one wave of two per SM: the second wave of a weight-gradient product cost as much as the first
                const bool wide = t.bn == 128 && K >= 4096;            // (these keep their 128-wide tiles, below)
                const long long tiles = (long long)((N + (wide ? 127 : 63)) / (wide ? 128 : 64)) * ((M + TCG_M - 1) / TCG_M) * Z;
                if (splits > 1 && tiles * splits > 296) splits = (int)std::max<long long>(1, 296 / tiles);
            }
            const int per = (n_chunks + splits - 1) / splits;
            splits = (n_chunks + per - 1) / per;                       // no empty split
            if ((long long)Z * splits * M * N > 2LL * TCG_MAX_SPLITS * (L.H + 1) * L.H) splits = 1;
            // 64-wide tiles when 128-wide ones would leave most of the 148 SMs without a CTA
            // (two CTAs fit one SM: 115 registers, <= 66 KB shared memory)
            if (t.bn == 128 && (long long)((N + 127) / 128) * ((M + TCG_M - 1) / TCG_M) * Z * splits < 296 && K < 4096) t.bn = 64;   // (contractions over a batch >= 4096 keep 128: half the re-reads of the A tile; measured 5 030 -> 5 240 updates/s at batch 4096, 6 470 -> 6 260 at batch 1024)
            t.tmem_cols = t.bn < 32 ? 32 : t.bn;
            t.splits = splits;
            t.partial = ws + L.o_part + (long long)(branch > 0 ? branch - 1 : 0) * 2LL * TCG_MAX_SPLITS * (L.H + 1) * L.H;
            const size_t stage = (size_t)TCG_M * TCG_KC * 2 + (size_t)t.bn * TCG_KC * 2;
            t.stages = std::max(2, std::min(t.bn <= 64 ? 4 : 3, (n_chunks + splits - 1) / splits));    // <= 96 KB: two CTAs per SM
            const size_t smem = std::max((size_t)t.stages * stage, (size_t)TCG_M * (t.bn + 1) * 4);
            const dim3 tgrid((N + t.bn - 1) / t.bn, (M + TCG_M - 1) / TCG_M, Z * splits);
            gemm_tc_kernel<<<tgrid, 256, smem, st>>>(t);
            if (splits > 1) splitk_reduce_kernel<<<dim3((unsigned)(((long long)M * N + 255) / 256), Z), 256, 0, st>>>(t);
            return;
        }
        const dim3 grid((N + 31) / 32, (M + 31) / 32, Z);
        const bool ak = a_cs == 1, bn = b_cs == 1;
        if (ak && bn) gemm_kernel<true, true><<<grid, 256, 0, st>>>(g);
        else if (ak) gemm_kernel<true, false><<<grid, 256, 0, st>>>(g);
        else if (bn) gemm_kernel<false, true><<<grid, 256, 0, st>>>(g);
        else gemm_kernel<false, false><<<grid, 256, 0, st>>>(g);
    }
    // forward of a 3-layer network: X [B][kin+1] -> h1, h2 [Z][B][H+1] (relu, ones column kept) -> out [Z][B][nout]
    // net_id: index of the (first) network's bf16 W1 image in the workspace (0 policy, 1 q1, 2 q2, 3 / 4 their targets);
    // -1 or the SIMT path: the three products as separate launches
    void forward(long long oX, int kin, long long oW, long long w_zs, const Net& n, int nout, long long oH1, long long oH2,
                 long long oOut, int Z, int net_id = -1, long long oEps = -1, long long oXa = 0, long long oLogp = 0, int keep = 1, bool img = true) const {
        const int B = L.B, H = L.H;
        const int ld = L.ldh;
        const long long hz = (long long)B * ld;
        if (L.tc && L.fused_fwd && net_id >= 0) {
            FwdTC f;
            f.X = ws + oX; f.kin1 = kin + 1; f.W = ws + oW; f.w_zs = w_zs; f.l0 = n.l0; f.l1 = n.l1; f.l2 = n.l2;
            f.w1p = reinterpret_cast<const unsigned char*>(ws + L.o_w1p) + (size_t)net_id * H * H * 2; f.w1p_zs = (long long)H * H * 2;
            // keep: 0 nothing (no backward through this pass), 1 activations + sign bits, 2 sign bits only (only an input gradient
            // goes back through it)
            f.H1 = keep == 1 ? ws + oH1 : nullptr; f.H2 = keep == 1 ? ws + oH2 : nullptr; f.Out = ws + oOut; f.hz = hz;
            const long long om = !keep ? -1 : (oH1 == L.o_hp1 ? L.o_mhp : (oH1 == L.o_hq1 ? L.o_mhq : -1));
            f.M1 = om >= 0 ? reinterpret_cast<uint32_t*>(ws + om) : nullptr;
            f.M2 = om >= 0 ? f.M1 + (long long)Z * B * 8 : nullptr;
            const long long rows = (long long)(B + TCG_M - 1) / TCG_M * TCG_M;
            const long long oi = (keep == 1 && img && L.dw_img) ? (oH1 == L.o_hp1 ? L.o_img_hp1 : (oH1 == L.o_hq1 ? L.o_img_hq1 : -1)) : -1;
            f.IMG1 = oi >= 0 ? reinterpret_cast<unsigned char*>(ws + oi) : nullptr; f.img_zs = rows * H * 2;
            f.IMG0 = oi >= 0 ? reinterpret_cast<unsigned char*>(ws + (oH1 == L.o_hp1 ? L.o_img_xp : L.o_img_xq)) : nullptr;
            f.B = B; f.H = H; f.ld = ld; f.nout = nout;
            f.eps = oEps >= 0 ? ws + oEps : nullptr; f.xa = ws + oXa; f.logp = ws + oLogp;
            f.tmem_cols = 32;
            while (f.tmem_cols < H) f.tmem_cols <<= 1;
            mlp_fwd_tc_kernel<<<dim3((B + TCG_M - 1) / TCG_M, Z), FUSED_NT, fwd_smem_bytes(H, kin + 1), st>>>(f);
            return;
        }
        gemm(oX, kin + 1, 1, 0, oW + n.l0, H, 1, w_zs, oH1, ld, hz, B, H, kin + 1, Z, EPI_RELU);
        gemm(oH1, ld, 1, hz, oW + n.l1, H, 1, w_zs, oH2, ld, hz, B, H, H + 1, Z, EPI_RELU);
        gemm(oH2, ld, 1, hz, oW + n.l2, nout, 1, w_zs, oOut, nout, (long long)B * nout, B, nout, H + 1, Z, EPI_STORE);
    }
    // all eight bf16 images of W1 for the fused forward / backward kernels (the critics' step of adam_kernel refreshes q1 / q2's)
    void pack_w1(cudaStream_t on) const {
        if (!(L.tc && L.fused_fwd)) return;
        const long long off[5] = {L.o_pol + L.pol.l1, L.o_q + L.q.l1, L.o_q + L.P_q + L.q.l1, L.o_qt + L.q.l1, L.o_qt + L.P_q + L.q.l1};
        PackNets pn;
        int n = 0;
        auto add = [&](int net, int img, int by_k) { pn.w_off[n] = off[net]; pn.img[n] = img; pn.by_k[n] = by_k; ++n; };
        add(0, 0, 0); add(3, 3, 0); add(4, 4, 0); add(0, 5, 1);
        add(1, 1, 0); add(2, 2, 0); add(1, 6, 1); add(2, 7, 1);
        for (int q = n; q < SAC_W1_IMAGES; ++q) { pn.w_off[q] = 0; pn.img[q] = 0; pn.by_k[q] = 0; }
        const int H = L.H;
        pack_w1_nets_kernel<<<dim3((H * (H >> 3) + 255) / 256, n), 256, 0, on>>>(ws, reinterpret_cast<unsigned char*>(ws + L.o_w1p), pn, H);
    }
    // backward: dOut [Z][B][nout] -> (optionally) weight gradients into the grad block, and d/dX of the first `dx_cols`
    // input columns starting at column dx_first (0 columns: stop at layer 0's weight gradient)
    void backward(long long oX, int kin, long long oW, long long w_zs, const Net& n, int nout, long long oH1, long long oH2,
                  long long oDOut, long long oDH2, long long oDH1, int Z, bool want_dw, long long oG, long long oDX, int dx_first,
                  int dx_cols, int bw_img = -1, int seed = 0) const {
        const int B = L.B, H = L.H;
        const int ld = L.ldh;
        const long long hz = (long long)B * ld, dz = (long long)B * H, oz = (long long)B * nout;
        if (L.tc && L.fused_fwd && bw_img >= 0 && dx_cols <= 1) {
            // input-gradient chain as ONE kernel; the three weight-gradient products beside / after it on the side streams
            const bool par2 = want_dw && side != nullptr;
            const bool seed_side = seed != 0 && side != nullptr;      // the loss kernel beside the backward kernel, which seeds itself
            if (par2 || seed_side) { cudaEventRecord(ev[0], st); cudaStreamWaitEvent(side, ev[0], 0); }             // dOut (or its inputs) ready
            if (seed == 1) critic_loss_kernel<<<1, 1024, 0, seed_side ? side : st>>>(ws, L, *hp);
            if (seed == 2) policy_seed_kernel<<<1, 1024, 0, seed_side ? side : st>>>(ws, L);
            if (want_dw) { branch = par2 ? 1 : 0; gemm(oH2, 1, ld, hz, oDOut, nout, 1, oz, oG + n.l2, nout, w_zs, H + 1, nout, B, Z, EPI_STORE); branch = 0; }
            BwdTC f;
            f.seed = seed_side ? seed : 0; f.oq = ws + L.o_oq; f.oqt = ws + L.o_oqt; f.logp_next = ws + L.o_logp_next; f.msk = ws + L.o_mask;
            f.rew = ws + L.o_r; f.alpha = ws + L.o_alpha; f.gamma = hp ? hp->gamma : 0.f;
            f.DOut = ws + oDOut; f.nout = nout; f.W = ws + oW; f.w_zs = w_zs; f.l0 = n.l0; f.l2 = n.l2;
            f.w1p = reinterpret_cast<const unsigned char*>(ws + L.o_w1p) + (size_t)bw_img * H * H * 2; f.w1p_zs = (long long)H * H * 2;
            f.M1 = reinterpret_cast<const uint32_t*>(ws + (oH1 == L.o_hp1 ? L.o_mhp : L.o_mhq)); f.M2 = f.M1 + (long long)Z * B * 8;
            f.DX = dx_cols > 0 ? ws + oDX : nullptr; f.dx_row = dx_first;
            const long long rows = (long long)(B + TCG_M - 1) / TCG_M * TCG_M;
            const int n_chunks = (int)(rows / 64);
            const bool dwi = want_dw && L.dw_img;
            const bool is_pol = oH1 == L.o_hp1;
            f.IMG2 = dwi ? reinterpret_cast<unsigned char*>(ws + (is_pol ? L.o_img_dhp2 : L.o_img_dhq2)) : nullptr; f.img_zs = rows * H * 2;
            f.IMG1 = dwi ? reinterpret_cast<unsigned char*>(ws + (is_pol ? L.o_img_dhp1 : L.o_img_dhq1)) : nullptr;
            // (with the images nobody reads the fp32 gradients: they are not stored)
            f.DH2 = (want_dw && !dwi) ? ws + oDH2 : nullptr; f.DH1 = (want_dw && !dwi) ? ws + oDH1 : nullptr;
            f.colsum = ws + L.o_colsum;
            f.B = B; f.H = H; f.tmem_cols = 32;
            while (f.tmem_cols < H) f.tmem_cols <<= 1;
            mlp_bwd_tc_kernel<<<dim3((B + TCG_M - 1) / TCG_M, Z), FUSED_NT, bwd_smem_bytes(H), st>>>(f);
            if (want_dw) {
                const bool two = par2 && side_b != nullptr;
                if (par2) { cudaEventRecord(ev[1], st); cudaStreamWaitEvent(side, ev[1], 0); if (two) cudaStreamWaitEvent(side_b, ev[1], 0); }
                if (dwi) {      // layer 1's weight gradient from the operand images the fused kernels wrote
                    DwImg t;
                    t.A = reinterpret_cast<const unsigned char*>(ws + (is_pol ? L.o_img_hp1 : L.o_img_hq1)); t.Bm = f.IMG2;
                    t.a_zs = t.b_zs = rows * H * 2; t.n_chunks = n_chunks; t.splits = (n_chunks + DWI_STAGES - 1) / DWI_STAGES; t.H = H; t.M = H;
                    t.partial = ws + L.o_part + 1LL * 2LL * TCG_MAX_SPLITS * (L.H + 1) * L.H;          // region of branch 2
                    cudaStream_t ss = par2 ? side : st;
                    dw_img_tc_kernel<<<dim3(H / 128, H / 128, Z * t.splits), 256, DWI_STAGES * 32768, ss>>>(t);
                    dw1_reduce_kernel<<<dim3((unsigned)(((long long)(H + 1) * H + 255) / 256), Z), 256, 0, ss>>>(
                        t, ws + oG + n.l1, w_zs, ws + L.o_colsum, (int)(rows / TCG_M));
                } else {
                    branch = par2 ? 2 : 0;
                    gemm(oH1, 1, ld, hz, oDH2, H, 1, dz, oG + n.l1, H, w_zs, H + 1, H, B, Z, EPI_STORE);
                }
                if (dwi) {      // layer 0's: the input image as the one row tile
                    DwImg t;
                    t.A = reinterpret_cast<const unsigned char*>(ws + (is_pol ? L.o_img_xp : L.o_img_xq)); t.Bm = f.IMG1;
                    t.a_zs = 0; t.b_zs = rows * H * 2; t.n_chunks = n_chunks; t.splits = (n_chunks + DWI_STAGES - 1) / DWI_STAGES; t.H = H; t.M = kin + 1;
                    t.partial = ws + L.o_part + 2LL * 2LL * TCG_MAX_SPLITS * (L.H + 1) * L.H;          // region of branch 3
                    cudaStream_t ss = two ? side_b : (par2 ? side : st);
                    dw_img_tc_kernel<<<dim3(H / 128, 1, Z * t.splits), 256, DWI_STAGES * 32768, ss>>>(t);
                    dw1_reduce_kernel<<<dim3((unsigned)(((long long)(kin + 1) * H + 255) / 256), Z), 256, 0, ss>>>(t, ws + oG + n.l0, w_zs, nullptr, 0);
                } else {
                    branch = par2 ? 3 : 0;
                    gemm(oX, 1, kin + 1, 0, oDH1, H, 1, dz, oG + n.l0, H, w_zs, kin + 1, H, B, Z, EPI_STORE);
                }
                branch = 0;
                if (two) { cudaEventRecord(ev_b[0], side_b); cudaStreamWaitEvent(st, ev_b[0], 0); }
                if (par2) { cudaEventRecord(ev[3], side); cudaStreamWaitEvent(st, ev[3], 0); }
            }
            return;
        }
        // The weight-gradient products (dW2, dW1, dW0) hang off the input-gradient chain dOut -> dH2 -> dH1 from the side: with a
        // side stream they run beside it (each waits for the one gradient it reads), which matters because every product of
        // the update leaves most of the GPU idle.  Without a side stream: in program order.
        const bool par = want_dw && side != nullptr;
        auto fork = [&](int e) { if (par) { cudaEventRecord(ev[e], st); cudaStreamWaitEvent(side, ev[e], 0); } };
        auto on_side = [&](int b) { branch = par ? b : 0; };
        // layer 2
        fork(0);                                                        // dOut is ready
        if (want_dw) { on_side(1); gemm(oH2, 1, ld, hz, oDOut, nout, 1, oz, oG + n.l2, nout, w_zs, H + 1, nout, B, Z, EPI_STORE); branch = 0; }
        gemm(oDOut, nout, 1, oz, oW + n.l2, 1, nout, w_zs, oDH2, H, dz, B, H, nout, Z, EPI_MASK, oH2, ld, hz);
        // layer 1
        fork(1);                                                        // dH2 is ready
        if (want_dw) { on_side(2); gemm(oH1, 1, ld, hz, oDH2, H, 1, dz, oG + n.l1, H, w_zs, H + 1, H, B, Z, EPI_STORE); branch = 0; }
        gemm(oDH2, H, 1, dz, oW + n.l1, 1, H, w_zs, oDH1, H, dz, B, H, H, Z, EPI_MASK, oH1, ld, hz);
        // layer 0
        fork(2);                                                        // dH1 is ready
        if (want_dw) { on_side(3); gemm(oX, 1, kin + 1, 0, oDH1, H, 1, dz, oG + n.l0, H, w_zs, kin + 1, H, B, Z, EPI_STORE); branch = 0; }
        if (dx_cols > 0)
            gemm(oDH1, H, 1, dz, oW + n.l0 + (long long)dx_first * H, 1, H, w_zs, oDX, dx_cols, (long long)B * dx_cols, B, dx_cols, H, Z,
                 EPI_STORE);
        if (par) { cudaEventRecord(ev[3], side); cudaStreamWaitEvent(st, ev[3], 0); }      // join: the gradients are complete
    }
};

// one update_parameters (main_ast.py:354): 37 kernels on `st` (43 on the tensor-core path: 6 split-k reduces)
static void enqueue_update(SitSac_* h, float* ws, const float* replay, long long capacity, const unsigned long long* n_rows,
                           float* losses, cudaStream_t st) {
    const Layout& L = h->L;
    const Hyper& hp = h->hp;
    const int B = L.B, H = L.H;
    const int eb = (B + 255) / 256;
    const bool par = st != nullptr && h->side != nullptr;   // branches on internal streams (all captured into the one graph)
    const Launcher X{st, ws, L, par ? h->side : nullptr, h->ev, 0, &hp, (par && L.tc && L.fused_fwd) ? h->side3 : nullptr, h->ev3};   // side_b only where the fused backward joins it
    const Launcher X2{par ? h->side2 : st, ws, L, nullptr, nullptr, 0};
    cudaStream_t s2 = X2.st;
    auto fork2 = [&](int e) { if (par) { cudaEventRecord(h->ev2[e], st); cudaStreamWaitEvent(s2, h->ev2[e], 0); } };
    auto join2 = [&](int e) { if (par) { cudaEventRecord(h->ev2[e], s2); cudaStreamWaitEvent(st, h->ev2[e], 0); } };
    // bf16 images of W1 from the weights as the caller left them (also after a load through sit_sac_field), beside the batch kernel
    if (par) { cudaEventRecord(h->ev3[2], st); cudaStreamWaitEvent(s2, h->ev3[2], 0); }
    X.pack_w1(s2);
    if (par) cudaEventRecord(h->ev3[3], s2);
    const int bt = L.tc ? 64 : 256;    // large batches: more, smaller CTAs (a row costs one dependent gather)
    batch_kernel<<<(B + bt - 1) / bt, bt, 0, st>>>(ws, L, hp, replay, capacity, n_rows);
    if (par) cudaStreamWaitEvent(st, h->ev3[3], 0);
    // critics on (s, a): needs the batch only -> beside the target computation
    fork2(0);
    X2.forward(L.o_xsa, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_oq, 2, 1);
    // target: a' ~ pi(s'), Q targets on (s', a')
    const bool fused = L.tc && L.fused_fwd;      // the policy's forward kernel then also applies the tanh-Gaussian head
    X.forward(L.o_xs2, S, L.o_pol, 0, L.pol, 2, L.o_hp1, L.o_hp2, L.o_op, 1, 0, L.o_eps_next, L.o_xs2a, L.o_logp_next, 0);   // a' ~ pi(s'): no backward, activations not kept
    if (!fused) head_kernel<<<eb, 256, 0, st>>>(ws, L, L.o_eps_next, L.o_xs2a, L.o_logp_next);
    X.forward(L.o_xs2a, SA, L.o_qt, L.P_q, L.q, 1, L.o_hqt1, L.o_hqt2, L.o_oqt, 2, 3, -1, 0, 0, 0);   // targets: no backward
    join2(1);
    // loss, backward, Adam (+ soft target update)
    if (!fused) critic_loss_kernel<<<1, 1024, 0, st>>>(ws, L, hp);      // (fused: beside the backward kernel, which seeds itself)
    // policy on s (its own buffers, the policy's weights): beside the critics' backward and Adam
    fork2(2);
    X2.forward(L.o_xs, S, L.o_pol, 0, L.pol, 2, L.o_hp1, L.o_hp2, L.o_op, 1, 0, L.o_eps_cur, L.o_xspi, L.o_logp_cur);
    if (!fused) head_kernel<<<eb, 256, 0, s2>>>(ws, L, L.o_eps_cur, L.o_xspi, L.o_logp_cur);
    X.backward(L.o_xsa, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_doq, L.o_dhq2, L.o_dhq1, 2, true, L.o_grad + L.P_pol, 0, 0, 0, 6, fused ? 1 : 0);
    adam_kernel<<<(unsigned)((2 * L.P_q + 255) / 256), 256, 0, st>>>(ws, L, hp, L.P_pol, 2 * L.P_q, L.o_qt - L.o_q, 1, (L.tc && L.fused_fwd) ? 1 : 0);   // + the updated critics' W1 images
    join2(3);
    // critics (updated) on (s, a~), d/da through the critics, policy backward, Adam
    X.forward(L.o_xspi, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_oq, 2, 1, -1, 0, 0, 2, false);   // only d Q / d a goes back through this pass: sign bits, no activations, no images
    if (!fused) policy_seed_kernel<<<1, 1024, 0, st>>>(ws, L);
    X.backward(L.o_xspi, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_doq, L.o_dhq2, L.o_dhq1, 2, false, 0, L.o_da, S, 1, 6, fused ? 2 : 0);
    policy_grad_kernel<<<eb, 256, 0, st>>>(ws, L);
    X.backward(L.o_xs, S, L.o_pol, 0, L.pol, 2, L.o_hp1, L.o_hp2, L.o_dop, L.o_dhp2, L.o_dhp1, 1, true, L.o_grad, 0, 0, 0, 5);
    adam_kernel<<<(unsigned)((L.P_pol + 255) / 256), 256, 0, st>>>(ws, L, hp, 0, L.P_pol, 0, 0);
    alpha_kernel<<<1, 32, 0, st>>>(ws, L, hp, losses);
    (void)H;
}

}  // namespace sac

extern "C" {

int sit_sac_create(const SitSacConfig* cfg, int device, SitSac* out) {
    if (!cfg || !out) return fail(SIT_E_INVALID, "null argument");
    if (cfg->hidden < 32 || cfg->hidden > 256 || cfg->hidden % 32) return fail(SIT_E_INVALID, "hidden must be a multiple of 32 in [32, 256]");
    if (cfg->batch_size < 1 || cfg->batch_size > 65536) return fail(SIT_E_INVALID, "batch_size must be in [1, 65536]");
    if (!(cfg->gamma >= 0.f && cfg->gamma <= 1.f) || !(cfg->tau >= 0.f && cfg->tau <= 1.f) || !(cfg->lr > 0.f) || !(cfg->alpha >= 0.f))
        return fail(SIT_E_INVALID, "bad gamma / tau / lr / alpha");
    if (cfg->target_update_interval < 1) return fail(SIT_E_INVALID, "target_update_interval must be >= 1");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(SIT_E_NODEVICE, "no CUDA device: there is no CPU fallback"); }
    if (device < 0 || device >= ndev) return fail(SIT_E_INVALID, "bad device index");
    SitSac_* h = new (std::nothrow) SitSac_();
    if (!h) return fail(SIT_E_INVALID, "out of memory");
    h->cfg = *cfg; h->device = device; h->graph = nullptr; h->side = nullptr; h->side2 = nullptr; h->side3 = nullptr;
    for (auto& e : h->ev) e = nullptr;
    for (auto& e : h->ev2) e = nullptr;
    for (auto& e : h->ev3) e = nullptr;
    // tensor cores from batch 1024 on: below that an update is 37 dependent few-microsecond launches and latency, not math,
    // is the cost (the SIMT products run at batch 64 in 220 us per update)
    const int tc = (cfg->tensor_cores && (cfg->batch_size >= 1024 || cfg->tensor_cores >= 2) && cfg->hidden >= 128) ? 1 : 0;
    sac::make_layout(h->L, cfg->hidden, cfg->batch_size, tc);
    h->hp.gamma = cfg->gamma; h->hp.tau = cfg->tau; h->hp.lr = cfg->lr; h->hp.alpha0 = cfg->alpha;
    h->hp.auto_alpha = cfg->automatic_entropy_tuning ? 1 : 0; h->hp.target_interval = cfg->target_update_interval;
    h->hp.seed = cfg->seed;
    {
        cudaError_t ce = cudaSetDevice(device);
        if (ce == cudaSuccess && tc) ce = cudaFuncSetAttribute(sac::gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (ce == cudaSuccess && tc)
            ce = cudaFuncSetAttribute(sac::mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sac::fwd_smem_bytes(cfg->hidden, sac::SA + 1));
        // two internal streams: independent parts of an update run as parallel branches of its CUDA graph (at batch 64 the
        // update is a chain of few-microsecond launches, at batch 4096 every product leaves most of the GPU idle)
        if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking);
        for (auto& e : h->ev) if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&h->side2, cudaStreamNonBlocking);
        for (auto& e : h->ev2) if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&h->side3, cudaStreamNonBlocking);
        for (auto& e : h->ev3) if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        if (ce == cudaSuccess && tc) ce = cudaFuncSetAttribute(sac::dw_img_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sac::DWI_STAGES * 32768);
        if (ce == cudaSuccess && tc) ce = cudaFuncSetAttribute(sac::mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sac::bwd_smem_bytes(cfg->hidden));
        if (ce != cudaSuccess) {
            if (h->side) cudaStreamDestroy(h->side);
            if (h->side2) cudaStreamDestroy(h->side2);
            if (h->side3) cudaStreamDestroy(h->side3);
            for (auto& e : h->ev) if (e) cudaEventDestroy(e);
            for (auto& e : h->ev2) if (e) cudaEventDestroy(e);
            for (auto& e : h->ev3) if (e) cudaEventDestroy(e);
            delete h;
            return fail(SIT_E_CUDA, std::string("learner set-up: ") + cudaGetErrorString(ce));
        }
    }
    *out = h;
    return SIT_OK;
}

int sit_sac_destroy(SitSac h) {
    if (!h) return SIT_OK;
    if (h->graph) cudaGraphExecDestroy(h->graph);
    if (h->side) cudaStreamDestroy(h->side);
    if (h->side2) cudaStreamDestroy(h->side2);
    if (h->side3) cudaStreamDestroy(h->side3);
    for (auto& e : h->ev) if (e) cudaEventDestroy(e);
    for (auto& e : h->ev2) if (e) cudaEventDestroy(e);
    for (auto& e : h->ev3) if (e) cudaEventDestroy(e);
    delete h;
    return SIT_OK;
}

int sit_sac_workspace_bytes(SitSac h, size_t* nbytes) {
    if (!h || !nbytes) return fail(SIT_E_INVALID, "null argument");
    *nbytes = (size_t)h->L.total * sizeof(float);
    return SIT_OK;
}

int sit_sac_field(SitSac h, const char* name, int64_t* offset, int64_t* count) {
    if (!h || !name || !offset || !count) return fail(SIT_E_INVALID, "null argument");
    const sac::Layout& L = h->L;
    struct { const char* n; long long o, c; } tab[] = {
        {"policy", L.o_pol, L.P_pol}, {"q1", L.o_q, L.P_q}, {"q2", L.o_q + L.P_q, L.P_q},
        {"q1_target", L.o_qt, L.P_q}, {"q2_target", L.o_qt + L.P_q, L.P_q}, {"log_alpha", L.o_log_alpha, 1},
        {"alpha", L.o_alpha, 1}, {"grad", L.o_grad, L.n_train}, {"adam_m", L.o_m, L.n_train}, {"adam_v", L.o_v, L.n_train},
        {"idx", L.o_idx, L.B}, {"eps_next", L.o_eps_next, L.B}, {"eps_cur", L.o_eps_cur, L.B},
        {"losses", L.o_losses, SIT_SAC_N_LOSS}, {"updates", L.o_updates, 2}, {"obs_scale", L.o_obs_scale, sac::S}};
    for (const auto& e : tab)
        if (std::strcmp(e.n, name) == 0) { *offset = e.o * (long long)sizeof(float); *count = e.c; return SIT_OK; }
    return fail(SIT_E_INVALID, std::string("unknown learner field: ") + name);
}

int sit_sac_init(SitSac h, void* workspace, void* stream) {
    if (!h || !workspace) return fail(SIT_E_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* ws = static_cast<float*>(workspace);
    const sac::Layout& L = h->L;
    const int H = L.H, B = L.B;
    CUDA_TRY(cudaMemsetAsync(ws, 0, (size_t)L.total * sizeof(float), st));
    auto init = [&](long long o, int K, int N, uint32_t id) {
        const long long n = (long long)(K + 1) * N;
        sac::init_layer_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ws + o, K, N, h->hp.seed, id);
    };
    init(L.o_pol + L.pol.l0, sac::S, H, 0); init(L.o_pol + L.pol.l1, H, H, 1); init(L.o_pol + L.pol.l2, H, 2, 2);
    for (int z = 0; z < 2; ++z) {
        const long long o = L.o_q + z * L.P_q;
        init(o + L.q.l0, sac::SA, H, 3 + 3 * z); init(o + L.q.l1, H, H, 4 + 3 * z); init(o + L.q.l2, H, 1, 5 + 3 * z);
    }
    CUDA_TRY(cudaMemcpyAsync(ws + L.o_qt, ws + L.o_q, (size_t)(2 * L.P_q) * sizeof(float), cudaMemcpyDeviceToDevice, st));
    sac::fill_kernel<<<1, 32, 0, st>>>(ws + L.o_alpha, 1, h->hp.alpha0);
    sac::ObsScale osc;
    for (int c = 0; c < sac::S; ++c) osc.v[c] = h->cfg.obs_scale[c];
    sac::obs_scale_kernel<<<1, 32, 0, st>>>(ws + L.o_obs_scale, osc);
    auto ones = [&](long long o, long long rows, int ld, int col = -1) {
        sac::ones_col_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(ws + o, rows, ld, col < 0 ? ld - 1 : col);
    };
    ones(L.o_xs, B, sac::S + 1); ones(L.o_xs2, B, sac::S + 1);
    ones(L.o_xsa, B, sac::SA + 1); ones(L.o_xs2a, B, sac::SA + 1); ones(L.o_xspi, B, sac::SA + 1);
    ones(L.o_hp1, B, L.ldh, H); ones(L.o_hp2, B, L.ldh, H); ones(L.o_hq1, 2LL * B, L.ldh, H); ones(L.o_hq2, 2LL * B, L.ldh, H);
    if (L.o_hqt1 != L.o_hq1) { ones(L.o_hqt1, 2LL * B, L.ldh, H); ones(L.o_hqt2, 2LL * B, L.ldh, H); }
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

int sit_sac_policy(SitSac h, void* workspace, SitPolicy* out) {
    if (!h || !workspace || !out) return fail(SIT_E_INVALID, "null argument");
    const sac::Layout& L = h->L;
    const float* p = static_cast<const float*>(workspace) + L.o_pol;
    out->mode = SIT_POLICY_GAUSSIAN; out->hidden = L.H;
    out->w0 = p + L.pol.l0; out->b0 = out->w0 + (long long)sac::S * L.H;
    out->w1 = p + L.pol.l1; out->b1 = out->w1 + (long long)L.H * L.H;
    out->w2 = p + L.pol.l2; out->b2 = out->w2 + (long long)L.H * 2;
    out->obs_scale = static_cast<const float*>(workspace) + L.o_obs_scale;
    return SIT_OK;
}

int sit_sac_update(SitSac h, void* workspace, const float* replay, int64_t capacity, const unsigned long long* n_rows,
                   int32_t n_updates, float* losses, int32_t use_graph, void* stream) {
    if (!h || !workspace || !replay) return fail(SIT_E_INVALID, "null argument");
    if (capacity < 1 || n_updates < 0) return fail(SIT_E_INVALID, "capacity must be >= 1 and n_updates >= 0");
    if (n_updates == 0) return SIT_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* ws = static_cast<float*>(workspace);
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (st != nullptr) CUDA_TRY(cudaStreamIsCapturing(st, &cs));
    if (!use_graph || st == nullptr || cs != cudaStreamCaptureStatusNone) {
        for (int i = 0; i < n_updates; ++i) sac::enqueue_update(h, ws, replay, capacity, n_rows, losses, st);
        CUDA_TRY(cudaGetLastError());
        return SIT_OK;
    }
    std::string key;
    key.append(reinterpret_cast<const char*>(&ws), sizeof(ws)); key.append(reinterpret_cast<const char*>(&replay), sizeof(replay));
    key.append(reinterpret_cast<const char*>(&capacity), sizeof(capacity)); key.append(reinterpret_cast<const char*>(&n_rows), sizeof(n_rows));
    key.append(reinterpret_cast<const char*>(&losses), sizeof(losses));
    if (!h->graph || key != h->graph_key) {
        if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; }
        CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        sac::enqueue_update(h, ws, replay, capacity, n_rows, losses, st);
        cudaGraph_t graph = nullptr;
        cudaError_t ce = cudaStreamEndCapture(st, &graph);
        if (ce != cudaSuccess) { if (graph) cudaGraphDestroy(graph); return fail(SIT_E_CUDA, std::string("graph capture: ") + cudaGetErrorString(ce)); }
        ce = cudaGraphInstantiate(&h->graph, graph, 0);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) { h->graph = nullptr; return fail(SIT_E_CUDA, std::string("graph instantiate: ") + cudaGetErrorString(ce)); }
        h->graph_key = key;
    }
    for (int i = 0; i < n_updates; ++i) CUDA_TRY(cudaGraphLaunch(h->graph, st));
    return SIT_OK;
}

}  // extern "C"
