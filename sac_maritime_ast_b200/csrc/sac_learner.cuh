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
// An update is 28 such products + 7 element-wise kernels; sizes are tiny (B = 64, H = 256: 0.16 GFLOP), so the cost is
// launch latency: the whole update is captured once into a CUDA graph and replayed.  All reductions have a fixed order:
// two runs with the same seed give bit-identical parameters.
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

// ---- workspace ------------------------------------------------------------------------------------------------------
struct Net { long long l0, l1, l2; };   // float offsets of the three augmented layer blocks inside a network block
struct Layout {
    int H, B;
    long long P_pol, P_q;               // floats per policy / critic block
    Net pol, q;
    // float offsets inside the workspace
    long long o_pol, o_q, o_qt, o_log_alpha, o_alpha, o_grad, o_m, o_v, n_train;   // n_train = P_pol + 2 P_q + 1 (log_alpha)
    long long o_xs, o_xs2, o_xsa, o_xs2a, o_xspi, o_r, o_mask;
    long long o_hp1, o_hp2, o_op, o_hq1, o_hq2, o_oq, o_oqt;
    long long o_doq, o_dhq2, o_dhq1, o_da, o_dop, o_dhp2, o_dhp1;
    long long o_eps_next, o_eps_cur, o_logp_next, o_logp_cur, o_idx, o_losses, o_updates, o_obs_scale, total;
};

static void make_layout(Layout& L, int H, int B) {
    L.H = H; L.B = B;
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
    L.o_hp1 = take(b * (H + 1)); L.o_hp2 = take(b * (H + 1)); L.o_op = take(b * 2);
    L.o_hq1 = take(2 * b * (H + 1)); L.o_hq2 = take(2 * b * (H + 1)); L.o_oq = take(2 * b); L.o_oqt = take(2 * b);
    L.o_doq = take(2 * b); L.o_dhq2 = take(2 * b * H); L.o_dhq1 = take(2 * b * H); L.o_da = take(2 * b);
    L.o_dop = take(b * 2); L.o_dhp2 = take(b * H); L.o_dhp1 = take(b * H);
    L.o_eps_next = take(b); L.o_eps_cur = take(b); L.o_logp_next = take(b); L.o_logp_cur = take(b);
    L.o_idx = take(b); L.o_losses = take(SIT_SAC_N_LOSS); L.o_updates = take(2);   // updates: one unsigned long long
    L.o_obs_scale = take(S);
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
__global__ void ones_col_kernel(float* p, long long rows, int ld) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < rows) p[i * ld + (ld - 1)] = 1.0f;
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
__global__ void adam_kernel(float* ws, const Layout L, const Hyper hp, long long first, long long count, long long target_delta,
                            int soft) {
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
};

namespace sac {

struct Launcher {
    cudaStream_t st;
    float* ws;
    const Layout& L;
    // C[z] (M x N, ldc) = epi(A[z] . B[z]);  A(m,k) = A[m a_rs + k a_cs], B(k,n) = B[k b_rs + n b_cs]
    void gemm(long long oA, int a_rs, int a_cs, long long a_zs, long long oB, int b_rs, int b_cs, long long b_zs, long long oC,
              int ldc, long long c_zs, int M, int N, int K, int Z, int epi, long long oH = 0, int ldh = 0, long long h_zs = 0) const {
        Gemm g;
        g.A = ws + oA; g.B = ws + oB; g.C = ws + oC; g.Hm = ws + oH;
        g.M = M; g.N = N; g.K = K; g.a_rs = a_rs; g.a_cs = a_cs; g.b_rs = b_rs; g.b_cs = b_cs; g.ldc = ldc; g.ldh = ldh;
        g.a_zs = a_zs; g.b_zs = b_zs; g.c_zs = c_zs; g.h_zs = h_zs; g.epi = epi;
        const dim3 grid((N + 31) / 32, (M + 31) / 32, Z);
        const bool ak = a_cs == 1, bn = b_cs == 1;
        if (ak && bn) gemm_kernel<true, true><<<grid, 256, 0, st>>>(g);
        else if (ak) gemm_kernel<true, false><<<grid, 256, 0, st>>>(g);
        else if (bn) gemm_kernel<false, true><<<grid, 256, 0, st>>>(g);
        else gemm_kernel<false, false><<<grid, 256, 0, st>>>(g);
    }
    // forward of a 3-layer network: X [B][kin+1] -> h1, h2 [Z][B][H+1] (relu, ones column kept) -> out [Z][B][nout]
    void forward(long long oX, int kin, long long oW, long long w_zs, const Net& n, int nout, long long oH1, long long oH2,
                 long long oOut, int Z) const {
        const int B = L.B, H = L.H;
        const long long hz = (long long)B * (H + 1);
        gemm(oX, kin + 1, 1, 0, oW + n.l0, H, 1, w_zs, oH1, H + 1, hz, B, H, kin + 1, Z, EPI_RELU);
        gemm(oH1, H + 1, 1, hz, oW + n.l1, H, 1, w_zs, oH2, H + 1, hz, B, H, H + 1, Z, EPI_RELU);
        gemm(oH2, H + 1, 1, hz, oW + n.l2, nout, 1, w_zs, oOut, nout, (long long)B * nout, B, nout, H + 1, Z, EPI_STORE);
    }
    // backward: dOut [Z][B][nout] -> (optionally) weight gradients into the grad block, and d/dX of the first `dx_cols`
    // input columns starting at column dx_first (0 columns: stop at layer 0's weight gradient)
    void backward(long long oX, int kin, long long oW, long long w_zs, const Net& n, int nout, long long oH1, long long oH2,
                  long long oDOut, long long oDH2, long long oDH1, int Z, bool want_dw, long long oG, long long oDX, int dx_first,
                  int dx_cols) const {
        const int B = L.B, H = L.H;
        const long long hz = (long long)B * (H + 1), dz = (long long)B * H, oz = (long long)B * nout;
        // layer 2
        if (want_dw) gemm(oH2, 1, H + 1, hz, oDOut, nout, 1, oz, oG + n.l2, nout, w_zs, H + 1, nout, B, Z, EPI_STORE);
        gemm(oDOut, nout, 1, oz, oW + n.l2, 1, nout, w_zs, oDH2, H, dz, B, H, nout, Z, EPI_MASK, oH2, H + 1, hz);
        // layer 1
        if (want_dw) gemm(oH1, 1, H + 1, hz, oDH2, H, 1, dz, oG + n.l1, H, w_zs, H + 1, H, B, Z, EPI_STORE);
        gemm(oDH2, H, 1, dz, oW + n.l1, 1, H, w_zs, oDH1, H, dz, B, H, H, Z, EPI_MASK, oH1, H + 1, hz);
        // layer 0
        if (want_dw) gemm(oX, 1, kin + 1, 0, oDH1, H, 1, dz, oG + n.l0, H, w_zs, kin + 1, H, B, Z, EPI_STORE);
        if (dx_cols > 0)
            gemm(oDH1, H, 1, dz, oW + n.l0 + (long long)dx_first * H, 1, H, w_zs, oDX, dx_cols, (long long)B * dx_cols, B, dx_cols, H, Z,
                 EPI_STORE);
    }
};

// one update_parameters (main_ast.py:354): ~35 kernels on `st`
static void enqueue_update(SitSac_* h, float* ws, const float* replay, long long capacity, const unsigned long long* n_rows,
                           float* losses, cudaStream_t st) {
    const Layout& L = h->L;
    const Hyper& hp = h->hp;
    const int B = L.B, H = L.H;
    const int eb = (B + 255) / 256;
    const Launcher X{st, ws, L};
    batch_kernel<<<eb, 256, 0, st>>>(ws, L, hp, replay, capacity, n_rows);
    // target: a' ~ pi(s'), Q targets on (s', a')
    X.forward(L.o_xs2, S, L.o_pol, 0, L.pol, 2, L.o_hp1, L.o_hp2, L.o_op, 1);
    head_kernel<<<eb, 256, 0, st>>>(ws, L, L.o_eps_next, L.o_xs2a, L.o_logp_next);
    X.forward(L.o_xs2a, SA, L.o_qt, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_oqt, 2);
    // critics on (s, a), loss, backward, Adam (+ soft target update)
    X.forward(L.o_xsa, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_oq, 2);
    critic_loss_kernel<<<1, 1024, 0, st>>>(ws, L, hp);
    X.backward(L.o_xsa, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_doq, L.o_dhq2, L.o_dhq1, 2, true, L.o_grad + L.P_pol, 0, 0, 0);
    adam_kernel<<<(unsigned)((2 * L.P_q + 255) / 256), 256, 0, st>>>(ws, L, hp, L.P_pol, 2 * L.P_q, L.o_qt - L.o_q, 1);
    // policy on s, critics (updated) on (s, a~), d/da through the critics, policy backward, Adam
    X.forward(L.o_xs, S, L.o_pol, 0, L.pol, 2, L.o_hp1, L.o_hp2, L.o_op, 1);
    head_kernel<<<eb, 256, 0, st>>>(ws, L, L.o_eps_cur, L.o_xspi, L.o_logp_cur);
    X.forward(L.o_xspi, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_oq, 2);
    policy_seed_kernel<<<1, 1024, 0, st>>>(ws, L);
    X.backward(L.o_xspi, SA, L.o_q, L.P_q, L.q, 1, L.o_hq1, L.o_hq2, L.o_doq, L.o_dhq2, L.o_dhq1, 2, false, 0, L.o_da, S, 1);
    policy_grad_kernel<<<eb, 256, 0, st>>>(ws, L);
    X.backward(L.o_xs, S, L.o_pol, 0, L.pol, 2, L.o_hp1, L.o_hp2, L.o_dop, L.o_dhp2, L.o_dhp1, 1, true, L.o_grad, 0, 0, 0);
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
    h->cfg = *cfg; h->device = device; h->graph = nullptr;
    sac::make_layout(h->L, cfg->hidden, cfg->batch_size);
    h->hp.gamma = cfg->gamma; h->hp.tau = cfg->tau; h->hp.lr = cfg->lr; h->hp.alpha0 = cfg->alpha;
    h->hp.auto_alpha = cfg->automatic_entropy_tuning ? 1 : 0; h->hp.target_interval = cfg->target_update_interval;
    h->hp.seed = cfg->seed;
    *out = h;
    return SIT_OK;
}

int sit_sac_destroy(SitSac h) {
    if (!h) return SIT_OK;
    if (h->graph) cudaGraphExecDestroy(h->graph);
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
    auto ones = [&](long long o, long long rows, int ld) { sac::ones_col_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(ws + o, rows, ld); };
    ones(L.o_xs, B, sac::S + 1); ones(L.o_xs2, B, sac::S + 1);
    ones(L.o_xsa, B, sac::SA + 1); ones(L.o_xs2a, B, sac::SA + 1); ones(L.o_xspi, B, sac::SA + 1);
    ones(L.o_hp1, B, H + 1); ones(L.o_hp2, B, H + 1); ones(L.o_hq1, 2LL * B, H + 1); ones(L.o_hq2, 2LL * B, H + 1);
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
