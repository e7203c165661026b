// (a7, a9, a10) per-sector factorisations, batched over all charge sectors of a tensor.
//
//  * SVD  : one-sided Jacobi (Hestenes) on the ROWS of a row-major working copy W (r x c, r <= c;
//           the transposed block is used when m > n so rows stay contiguous/coalesced).  Rotations
//           are accumulated in Z (r x r, unitary):  W_final = Z W_0  =>  W_0 = Z^H diag(S) Vhat.
//           Round-robin (tournament) ordering gives r/2 disjoint row pairs per round, one warp per
//           pair; a "team" of G CTAs (a thread-block cluster, hardware barrier between rounds)
//           works on one matrix, teams work on different sectors concurrently.
//           Replaces block_backend->matrix_svd per sector (abelian.cpp:3517; numpy.cpp:1247-1297,
//           scipy.linalg.svd gesdd).  Singular values come out with Jacobi's high relative accuracy.
//  * EIGH : the same kernel on W_0 = H + sigma*I (sigma = 1.5*||H||_F, positive definite), whose
//           left/right singular vectors coincide with the eigenvectors; lambda = S - sigma.  The eigenvectors are
//           read off the normalised rows of W_final (W_0^H W_0 = Vhat^H S^2 Vhat), so the left factor Z is neither
//           initialised nor updated: half of the row updates of a sweep.
//           Replaces block_backend->eigh (abelian.cpp:1781; numpy.cpp:658-680 np.linalg.eigh).
//  * QR   : blocked Householder QR (compact WY, panels of QR_NB reflectors) on the transposed copy (columns of
//           A become contiguous rows): the panel is factored in shared memory by CTA 0 of the team, every
//           trailing column takes the whole panel in one pass (one warp per column), then blocked backward
//           accumulation of Q.
//           Replaces block_backend->matrix_qr(block, full=false) (abelian.cpp:3123;
//           numpy.cpp:1236-1245 scipy.linalg.qr economic).
//  Gauge: singular/eigen vectors are defined up to phases (and rotations in degenerate spaces), R's
//  diagonal up to phases; parity with the reference is on S / W values, reconstruction and isometry.
#include "common.cuh"
#include <cstdio>
#include <vector>
#include <cmath>
#include <algorithm>
#include <cstdlib>

namespace cyb {

constexpr int FZ_THREADS = 1024;
constexpr int FZ_WARPS = FZ_THREADS / 32;
constexpr int FZ_MAX_SWEEPS = 60;
constexpr int QR_NB = 4;                // reflectors per panel of the blocked Householder QR
constexpr double FZ_EPS = 2.220446049250313e-16;
constexpr double FZ_TINY = 1e-280;

// ---- element traits ------------------------------------------------------------------------
template <bool C> struct El;
template <> struct El<false> {
    using T = double;
    static __device__ __forceinline__ T zero() { return 0.0; }
    static __device__ __forceinline__ T one() { return 1.0; }
    static __device__ __forceinline__ T conj(T x) { return x; }
    static __device__ __forceinline__ T mul(T a, T b) { return a * b; }
    static __device__ __forceinline__ T add(T a, T b) { return a + b; }
    static __device__ __forceinline__ T sub(T a, T b) { return a - b; }
    static __device__ __forceinline__ T scale(T a, double s) { return a * s; }
    static __device__ __forceinline__ double abs2(T a) { return a * a; }
    static __device__ __forceinline__ double re(T a) { return a; }
    static __device__ __forceinline__ T from_re(double a) { return a; }
    static __device__ __forceinline__ T ldcg(const T* p) { return __ldcg(p); }
    static __device__ __forceinline__ void stcg(T* p, T v) { __stcg(p, v); }
    static __device__ __forceinline__ T warp_sum(T v) { return cyb::warp_sum(v); }
};
template <> struct El<true> {
    using T = double2;
    static __device__ __forceinline__ T zero() { return make_double2(0.0, 0.0); }
    static __device__ __forceinline__ T one() { return make_double2(1.0, 0.0); }
    static __device__ __forceinline__ T conj(T x) { return make_double2(x.x, -x.y); }
    static __device__ __forceinline__ T mul(T a, T b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
    static __device__ __forceinline__ T add(T a, T b) { return make_double2(a.x + b.x, a.y + b.y); }
    static __device__ __forceinline__ T sub(T a, T b) { return make_double2(a.x - b.x, a.y - b.y); }
    static __device__ __forceinline__ T scale(T a, double s) { return make_double2(a.x * s, a.y * s); }
    static __device__ __forceinline__ double abs2(T a) { return a.x * a.x + a.y * a.y; }
    static __device__ __forceinline__ double re(T a) { return a.x; }
    static __device__ __forceinline__ T from_re(double a) { return make_double2(a, 0.0); }
    static __device__ __forceinline__ T ldcg(const T* p) { return __ldcg(p); }
    static __device__ __forceinline__ void stcg(T* p, T v) { __stcg(p, v); }
    static __device__ __forceinline__ T warp_sum(T v) { return make_double2(cyb::warp_sum(v.x), cyb::warp_sum(v.y)); }
};

// ---- teams: G CTAs of one thread-block cluster cooperate on one matrix ----------------------
template <int G> struct Team {
    static __device__ __forceinline__ unsigned rank()
    {
        if (G == 1) return 0u;
        unsigned r;
        asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
        return r;
    }
    static __device__ __forceinline__ unsigned id()
    {
        if (G == 1) return blockIdx.x;
        unsigned r;
        asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
        return r;
    }
    static __device__ __forceinline__ void sync()
    {
        if (G == 1) { __syncthreads(); }
        else {
            __threadfence();
            asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
            asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
        }
    }
};

struct FzDesc {
    int64_t m, n;                 // logical block shape
    int64_t a_off, o0_off, o1_off, o2_off;  // input and up to three outputs (element offsets)
    int64_t r, c;                 // working shape (r <= c for svd/eigh)
    int64_t w_off, z_off;         // element offsets (of T) into the workspace
    int64_t aux_off;              // double offsets into the aux workspace
    int64_t transposed;           // working copy is A^T
    double shift;                 // eigh: filled by the kernel
};

// ---- SVD / EIGH ------------------------------------------------------------------------------
// aux layout per matrix (doubles): coef[2r] (r elements of T, 16-byte aligned) | S_raw[r] | pos[r] (int64) |
//                                   rot_count[FZ_MAX_SWEEPS+1] (uint64) | pdiag[c] | misc[8]
__host__ __device__ inline int64_t svd_aux_doubles(int64_t r, int64_t c) { return 2 * r + 2 * r + (FZ_MAX_SWEEPS + 1) + c + 8; }

// PH selects the phases a launch runs: 1 = load the working copy, 2 = scalar Jacobi sweeps,
// 4 = ordering / normalisation / outputs.  Small sectors run all three in one launch (PH = 7); large
// real sectors run 1, then the block-Jacobi rounds (bj_pair_kernel below), then 4.
constexpr int PH_INIT = 1, PH_SWEEPS = 2, PH_FINISH = 4, PH_ALL = 7;

template <bool CX, int G, bool EIGH, int PH>
__global__ void __launch_bounds__(FZ_THREADS, 1)
jacobi_svd_kernel(const void* __restrict__ a_, void* __restrict__ o0_, double* __restrict__ s_out,
                  void* __restrict__ o2_, void* __restrict__ ws_, double* __restrict__ aux_,
                  FzDesc* __restrict__ descs, const int64_t* __restrict__ sel, int64_t* __restrict__ info)
{
    using E = El<CX>;
    using T = typename E::T;
    const int64_t mat = sel[Team<G>::id()];
    const FzDesc D = descs[mat];
    const int64_t r = D.r, c = D.c, m = D.m, n = D.n;
    const T* A = static_cast<const T*>(a_) + D.a_off;
    T* W = static_cast<T*>(ws_) + D.w_off;
    T* Z = static_cast<T*>(ws_) + D.z_off;
    T* coef = reinterpret_cast<T*>(aux_ + D.aux_off);   // r elements (<= 2r doubles)
    double* S_raw = aux_ + D.aux_off + 2 * r;
    int64_t* pos = reinterpret_cast<int64_t*>(S_raw + r);
    unsigned long long* rot_count = reinterpret_cast<unsigned long long*>(S_raw + 2 * r);
    double* pdiag = S_raw + 2 * r + (FZ_MAX_SWEEPS + 1);
    double* misc = pdiag + c;
    double sigma = 0.0;

    const unsigned trank = Team<G>::rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t gthreads = (int64_t)G * FZ_THREADS, gtid = (int64_t)trank * FZ_THREADS + tid;
    const int64_t gwarps = (int64_t)G * FZ_WARPS, gwarp = (int64_t)trank * FZ_WARPS + warp;
    __shared__ double s_red[FZ_WARPS];
    __shared__ unsigned long long s_cnt;

    if (!(PH & PH_INIT)) sigma = __ldcg(&misc[2]);
    // ---- load working copy, Z = I ----
    if (!(PH & PH_INIT)) {
    } else if (EIGH) {
        // sigma = 1.5 * ||H||_F  (every CTA of the team computes it redundantly, deterministic)
        double acc = 0.0;
        for (int64_t e = tid; e < r * c; e += FZ_THREADS) acc += E::abs2(A[e]);
        acc = warp_sum(acc);
        if (lane == 0) s_red[warp] = acc;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int w = 0; w < FZ_WARPS; ++w) t += s_red[w];
            s_red[0] = 1.5 * sqrt(t);
        }
        __syncthreads();
        sigma = s_red[0];
        for (int64_t e = gtid; e < r * c; e += gthreads) {
            const int64_t i = e / c, j = e % c;
            // symmetrise: use (H + H^H)/2 so that tiny asymmetries of the input do not matter
            T v = E::scale(E::add(A[i * c + j], E::conj(A[j * c + i])), 0.5);
            if (i == j) v = E::add(v, E::from_re(sigma));
            E::stcg(&W[e], v);
        }
    } else if (D.transposed) {
        for (int64_t e = gtid; e < r * c; e += gthreads) {
            const int64_t i = e / c, j = e % c;      // W[i][j] = A[j][i], A is m x n = c x r
            E::stcg(&W[e], A[j * n + i]);
        }
    } else {
        for (int64_t e = gtid; e < r * c; e += gthreads) E::stcg(&W[e], A[e]);
    }
    if (PH & PH_INIT) {
        if (!EIGH)
            for (int64_t e = gtid; e < r * r; e += gthreads) E::stcg(&Z[e], (e / r == e % r) ? E::one() : E::zero());
        for (int64_t e = gtid; e < FZ_MAX_SWEEPS + 1; e += gthreads) rot_count[e] = 0ull;
        if (gtid == 0) __stcg(&misc[2], sigma);
    }
    Team<G>::sync();
    if (!(PH & (PH_SWEEPS | PH_FINISH))) return;

    // ---- Jacobi sweeps ----
    const double tol = FZ_EPS * sqrt((double)c);
    const int64_t rp = r + (r & 1);          // players (padded to even)
    const int64_t rounds = rp - 1, npairs = rp / 2;
    int sweeps_done = 0;
    bool converged = (r < 2);
    for (int sweep = 0; (PH & PH_SWEEPS) && sweep < FZ_MAX_SWEEPS && !converged; ++sweep) {
        unsigned long long my_rot = 0ull;
        for (int64_t round = 0; round < rounds; ++round) {
            for (int64_t k = gwarp; k < npairs; k += gwarps) {
                int64_t i, j;
                if (k == 0) { i = rp - 1; j = round; }
                else { i = (round + k) % (rp - 1); j = (round - k + (rp - 1)) % (rp - 1); }
                if (i > j) { int64_t t = i; i = j; j = t; }
                if (j >= r) continue;        // dummy player
                T* x = W + i * c;
                T* y = W + j * c;
                double alpha = 0.0, beta = 0.0;
                T gamma = E::zero();
                for (int64_t q = lane; q < c; q += 32) {
                    const T xv = E::ldcg(x + q), yv = E::ldcg(y + q);
                    alpha += E::abs2(xv);
                    beta += E::abs2(yv);
                    gamma = E::add(gamma, E::mul(xv, E::conj(yv)));
                }
                alpha = warp_sum(alpha);
                beta = warp_sum(beta);
                gamma = E::warp_sum(gamma);
                const double g2 = E::abs2(gamma);
                if (alpha < FZ_TINY || beta < FZ_TINY) continue;
                if (g2 <= tol * tol * alpha * beta) continue;
                const double ga = sqrt(g2);
                // y' = phase * y with phase = gamma/|gamma| makes <x, y'> = |gamma| real
                const T phase = E::scale(gamma, 1.0 / ga);
                const double zeta = (beta - alpha) / (2.0 * ga);
                const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int64_t q = lane; q < c; q += 32) {
                    const T xv = E::ldcg(x + q), yv = E::mul(phase, E::ldcg(y + q));
                    E::stcg(x + q, E::sub(E::scale(xv, cs), E::scale(yv, sn)));
                    E::stcg(y + q, E::add(E::scale(xv, sn), E::scale(yv, cs)));
                }
                if (!EIGH) {
                    T* zx = Z + i * r;
                    T* zy = Z + j * r;
                    for (int64_t q = lane; q < r; q += 32) {
                        const T xv = E::ldcg(zx + q), yv = E::mul(phase, E::ldcg(zy + q));
                        E::stcg(zx + q, E::sub(E::scale(xv, cs), E::scale(yv, sn)));
                        E::stcg(zy + q, E::add(E::scale(xv, sn), E::scale(yv, cs)));
                    }
                }
                ++my_rot;
            }
            Team<G>::sync();
        }
        // team-wide rotation count of this sweep
        if (tid == 0) s_cnt = 0ull;
        __syncthreads();
        if (lane == 0 && my_rot) atomicAdd(&s_cnt, my_rot);
        __syncthreads();
        if (tid == 0 && s_cnt) atomicAdd(&rot_count[sweep], s_cnt);
        Team<G>::sync();
        const unsigned long long total = __ldcg(&rot_count[sweep]);
        sweeps_done = sweep + 1;
        converged = (total == 0ull);
    }

    if (!(PH & PH_FINISH)) return;
    // ---- singular values, ordering ----
    for (int64_t i = gwarp; i < r; i += gwarps) {
        double acc = 0.0;
        for (int64_t q = lane; q < c; q += 32) acc += E::abs2(E::ldcg(W + i * c + q));
        acc = warp_sum(acc);
        if (lane == 0) __stcg(&S_raw[i], sqrt(acc));
    }
    Team<G>::sync();
    // descending for svd, ascending (of S - sigma, i.e. of S) for eigh; stable rank sort
    for (int64_t i = gtid; i < r; i += gthreads) {
        const double si = __ldcg(&S_raw[i]);
        int64_t rank = 0;
        for (int64_t j = 0; j < r; ++j) {
            const double sj = __ldcg(&S_raw[j]);
            if (EIGH) rank += (sj < si) || (sj == si && j < i);
            else rank += (sj > si) || (sj == si && j < i);
        }
        __stcg(&pos[i], rank);
    }
    Team<G>::sync();

    // ---- normalise rows (Vhat); null rows are completed afterwards ----
    for (int64_t i = gwarp; i < r; i += gwarps) {
        const double s = __ldcg(&S_raw[i]);
        if (s * s < FZ_TINY) continue;
        const double inv = 1.0 / s;
        for (int64_t q = lane; q < c; q += 32) E::stcg(W + i * c + q, E::scale(E::ldcg(W + i * c + q), inv));
    }
    Team<G>::sync();

    if (!EIGH) {
        // completion of exactly-null rows by Gram-Schmidt against all finished rows (CTA 0 only)
        if (trank == 0) {
            int64_t n_null = 0;
            for (int64_t i = 0; i < r; ++i) { const double s = __ldcg(&S_raw[i]); n_null += (s * s < FZ_TINY); }
            if (n_null > 0) {
                // pdiag[j] = sum over valid rows |Vhat[i][j]|^2
                for (int64_t j = tid; j < c; j += FZ_THREADS) {
                    double acc = 0.0;
                    for (int64_t i = 0; i < r; ++i) {
                        const double s = __ldcg(&S_raw[i]);
                        if (s * s < FZ_TINY) continue;
                        acc += E::abs2(E::ldcg(W + i * c + j));
                    }
                    pdiag[j] = acc;
                }
                __syncthreads();
                for (int64_t i0 = 0; i0 < r; ++i0) {
                    const double s0 = __ldcg(&S_raw[i0]);
                    if (!(s0 * s0 < FZ_TINY)) continue;
                    // candidate column with the smallest projector diagonal
                    if (tid == 0) {
                        int64_t best = 0; double bv = pdiag[0];
                        for (int64_t j = 1; j < c; ++j) if (pdiag[j] < bv) { bv = pdiag[j]; best = j; }
                        misc[0] = (double)best;
                    }
                    __syncthreads();
                    const int64_t cand = (int64_t)misc[0];
                    T* v = W + i0 * c;
                    for (int64_t j = tid; j < c; j += FZ_THREADS) E::stcg(v + j, j == cand ? E::one() : E::zero());
                    __syncthreads();
                    for (int pass = 0; pass < 2; ++pass) {
                        // coef[k] = <row_k, v> = sum_j conj(row_k[j]) v[j]
                        for (int64_t k = warp; k < r; k += FZ_WARPS) {
                            const double sk = __ldcg(&S_raw[k]);
                            const bool valid = (k != i0) && (!(sk * sk < FZ_TINY) || k < i0);
                            T acc = E::zero();
                            if (valid)
                                for (int64_t j = lane; j < c; j += 32) acc = E::add(acc, E::mul(E::conj(E::ldcg(W + k * c + j)), E::ldcg(v + j)));
                            acc = E::warp_sum(acc);
                            if (lane == 0) coef[k] = acc;
                        }
                        __syncthreads();
                        for (int64_t j = tid; j < c; j += FZ_THREADS) {
                            T acc = E::ldcg(v + j);
                            for (int64_t k = 0; k < r; ++k) acc = E::sub(acc, E::mul(coef[k], E::ldcg(W + k * c + j)));
                            E::stcg(v + j, acc);
                        }
                        __syncthreads();
                    }
                    double acc = 0.0;
                    for (int64_t j = tid; j < c; j += FZ_THREADS) acc += E::abs2(E::ldcg(v + j));
                    acc = warp_sum(acc);
                    if (lane == 0) s_red[warp] = acc;
                    __syncthreads();
                    if (tid == 0) { double t = 0.0; for (int w = 0; w < FZ_WARPS; ++w) t += s_red[w]; misc[1] = t; }
                    __syncthreads();
                    const double inv = 1.0 / sqrt(misc[1]);
                    for (int64_t j = tid; j < c; j += FZ_THREADS) {
                        const T nv = E::scale(E::ldcg(v + j), inv);
                        E::stcg(v + j, nv);
                        pdiag[j] += E::abs2(nv);
                    }
                    __syncthreads();
                }
            }
        }
        Team<G>::sync();
    }

    // ---- outputs ----
    if (EIGH) {
        // w (ascending) and V[j][p] = conj(Vhat[i][j]) with p = pos[i] (Vhat = the normalised rows of W, c == r)
        T* V = static_cast<T*>(o2_) + D.o2_off;
        for (int64_t i = gtid; i < r; i += gthreads) s_out[D.o1_off + __ldcg(&pos[i])] = __ldcg(&S_raw[i]) - sigma;
        for (int64_t e = gtid; e < r * r; e += gthreads) {
            const int64_t j = e / r, i = e % r;   // consecutive threads -> consecutive i (scattered columns)
            V[j * r + __ldcg(&pos[i])] = E::conj(E::ldcg(W + i * c + j));
        }
    } else {
        T* U = static_cast<T*>(o0_) + D.o0_off;
        T* Vh = static_cast<T*>(o2_) + D.o2_off;
        for (int64_t i = gtid; i < r; i += gthreads) s_out[D.o1_off + __ldcg(&pos[i])] = __ldcg(&S_raw[i]);
        if (!D.transposed) {
            // U (m x k = r x r): U[a][p] = conj(Z[i][a]);  Vh (k x n = r x c): Vh[p][:] = Vhat[i][:]
            for (int64_t e = gtid; e < r * r; e += gthreads) {
                const int64_t a = e / r, i = e % r;
                U[a * r + __ldcg(&pos[i])] = E::conj(E::ldcg(Z + i * r + a));
            }
            for (int64_t e = gtid; e < r * c; e += gthreads) {
                const int64_t i = e / c, q = e % c;
                Vh[__ldcg(&pos[i]) * c + q] = E::ldcg(W + e);
            }
        } else {
            // A = W^T:  U (m x k = c x r): U[q][p] = Vhat[i][q];  Vh (k x n = r x r): Vh[p][b] = conj(Z[i][b])... see header
            for (int64_t e = gtid; e < r * c; e += gthreads) {
                const int64_t q = e / r, i = e % r;
                U[q * r + __ldcg(&pos[i])] = E::ldcg(W + i * c + q);
            }
            for (int64_t e = gtid; e < r * r; e += gthreads) {
                const int64_t i = e / r, b = e % r;
                Vh[__ldcg(&pos[i]) * r + b] = E::conj(E::ldcg(Z + e));
            }
        }
    }
    if ((PH & PH_SWEEPS) && gtid == 0) info[mat] = converged ? (sweeps_done > 0 ? sweeps_done : 1) : -(sweeps_done > 0 ? sweeps_done : 1);
    (void)m;
}

}  // namespace cyb
#include "block_jacobi.cuh"

namespace cyb {
// launch with (or without) programmatic stream serialization; the kernels call bj_pdl_enter() first thing either way
template <class... KArgs, class... Args>
static cudaError_t bj_launch(bool pdl, void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t stream, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(block, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
}  // namespace cyb

#include "qr_blocked.cuh"
namespace cyb {

// ---- QR ---------------------------------------------------------------------------------------
// Working copy Wt = A^T (n x m): row i of Wt is column i of A.  Hermitian Householder reflectors
// H_j = I - tau_j u_j u_j^H (tau real), u_j stored in Wt[j][j+1:m] with its head in uhead[j].
// BLK (large real sectors): two-level blocking.  An outer panel of QB_NBO = 32 columns is factored by CTA 0 of the team
// (inner panels of QR_NB reflectors, applied only inside the outer panel); the trailing matrix and, in the second phase,
// Q take the 32 reflectors at once as compact-WY GEMMs on the FP64 tensor pipe, split over the CTAs of the team
// (qr_blocked.cuh).
template <bool CX, int G, bool BLK>
__global__ void __launch_bounds__(BLK ? QB_THREADS : FZ_THREADS, 1)
householder_qr_kernel(const void* __restrict__ a_, void* __restrict__ q_, void* __restrict__ r_,
                      void* __restrict__ ws_, double* __restrict__ aux_, FzDesc* __restrict__ descs,
                      const int64_t* __restrict__ sel, int use_smem)
{
    using E = El<CX>;
    using T = typename E::T;
    extern __shared__ __align__(16) unsigned char qr_smem[];
    T* s_U = reinterpret_cast<T*>(qr_smem);      // use_smem: the QR_NB x m panel (working rows, then reflectors)
    const int64_t mat = sel[Team<G>::id()];
    const FzDesc D = descs[mat];
    const int64_t m = D.m, n = D.n, k = m < n ? m : n;
    const T* A = static_cast<const T*>(a_) + D.a_off;
    T* Wt = static_cast<T*>(ws_) + D.w_off;     // n x m
    T* Qt = static_cast<T*>(ws_) + D.z_off;     // k x m
    T* uhead = reinterpret_cast<T*>(aux_ + D.aux_off);   // k elements (16-byte aligned)
    double* tau = aux_ + D.aux_off + 2 * k;              // k
    T* Q = static_cast<T*>(q_) + D.o0_off;      // m x k
    T* R = static_cast<T*>(r_) + D.o1_off;      // k x n
    constexpr int NT = BLK ? QB_THREADS : FZ_THREADS, NW = NT / 32;   // threads / warps per CTA
    const unsigned trank = Team<G>::rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t gthreads = (int64_t)G * NT, gtid = (int64_t)trank * NT + tid;
    const int64_t gwarps = (int64_t)G * NW, gwarp = (int64_t)trank * NW + warp;
    // CYB_QR_PROFILE=1: wall time of the phases as seen by thread 0 of CTA 0 of every team (device printf at the end)
    const bool prof = (use_smem & 2) != 0;
    use_smem &= 1;
    unsigned long long pt[13] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, pl = 0;
    auto tick = [&](int slot) {
        if (!prof) return;
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (slot >= 0) pt[slot] += now - pl;
        pl = now;
    };
    tick(-1);

    for (int64_t e = gtid; e < n * m; e += gthreads) {
        const int64_t i = e / m, q = e % m;
        E::stcg(&Wt[e], A[q * n + i]);
    }
    Team<G>::sync();
    tick(0);

    // Blocked Householder (compact WY, panels of QR_NB reflectors): H_j0 ... H_j0+nb-1 = I - U T U^H.
    // The panel is factored by CTA 0 of the team (unblocked, on the nb panel rows only); every trailing row
    // then takes ALL nb reflectors in one pass (nb dots, w = T^H d, one update) -- a quarter of the L2
    // traffic and of the team barriers of the column-by-column version.
    __shared__ T s_T[QR_NB][QR_NB];
    __shared__ T s_z[QR_NB][QR_NB];
    // element q of reflector j (j zeros, head, then the stored tail)
    int64_t cur_j0 = 0;                           // panel currently held in s_U
    auto uget = [&](int64_t j, int64_t q) -> T {
        if (use_smem) return s_U[(j - cur_j0) * m + q];
        return q > j ? E::ldcg(Wt + j * m + q) : (q == j ? E::ldcg(&uhead[j]) : E::zero());
    };
    // reflectors of the panel [j0, j0 + nb) -> s_U (zeros before the head), every CTA of the team
    auto load_panel = [&](int64_t j0, int nb) {
        cur_j0 = j0;
        if (!use_smem) return;
        __syncthreads();
        for (int64_t e = tid; e < (int64_t)nb * m; e += NT) {
            const int64_t j = j0 + e / m, q = e % m;
            s_U[e] = q > j ? E::ldcg(Wt + j * m + q) : (q == j ? E::ldcg(&uhead[j]) : E::zero());
        }
        __syncthreads();
    };
    // T of the panel [j0, j0 + nb) into s_T (every CTA builds its own copy): z[b][a] = u_b^H u_a, b < a
    auto build_T = [&](int64_t j0, int nb) {
        for (int pr = warp; pr < nb * nb; pr += NW) {
            const int bb = pr / nb, aa = pr % nb;
            if (bb >= aa) continue;
            T dot = E::zero();
            for (int64_t q = j0 + aa + lane; q < m; q += 32) dot = E::add(dot, E::mul(E::conj(uget(j0 + bb, q)), uget(j0 + aa, q)));
            dot = E::warp_sum(dot);
            if (lane == 0) s_z[bb][aa] = dot;
        }
        __syncthreads();
        if (tid == 0) {
            for (int aa = 0; aa < nb; ++aa) {
                const double ta = __ldcg(&tau[j0 + aa]);
                for (int bb = 0; bb < QR_NB; ++bb) s_T[bb][aa] = E::zero();
                s_T[aa][aa] = E::from_re(ta);
                for (int bb = 0; bb < aa; ++bb) {          // T[0:a, a] = -tau_a T[0:a, 0:a] z[0:a, a]
                    T acc = E::zero();
                    for (int cc = bb; cc < aa; ++cc) acc = E::add(acc, E::mul(s_T[bb][cc], s_z[cc][aa]));
                    s_T[bb][aa] = E::scale(acc, -ta);
                }
            }
        }
        __syncthreads();
    };
    // y <- (I - U X U^H) y for one row y (length m) by one warp; X = T^H (ADJ) or T
    auto apply_panel = [&](T* y, int64_t j0, int nb, bool adj) {
        T d[QR_NB];
#pragma unroll
        for (int aa = 0; aa < QR_NB; ++aa) d[aa] = E::zero();
#pragma unroll 4
        for (int64_t q = j0 + lane; q < m; q += 32) {
            const T yv = E::ldcg(y + q);
#pragma unroll
            for (int aa = 0; aa < QR_NB; ++aa)
                if (aa < nb) d[aa] = E::add(d[aa], E::mul(E::conj(uget(j0 + aa, q)), yv));
        }
        T w[QR_NB];
#pragma unroll
        for (int aa = 0; aa < QR_NB; ++aa) d[aa] = E::warp_sum(d[aa]);
#pragma unroll
        for (int aa = 0; aa < QR_NB; ++aa) {
            T acc = E::zero();
#pragma unroll
            for (int bb = 0; bb < QR_NB; ++bb) {
                if (bb >= nb || aa >= nb) continue;
                if (adj) { if (bb <= aa) acc = E::add(acc, E::mul(E::conj(s_T[bb][aa]), d[bb])); }
                else { if (bb >= aa) acc = E::add(acc, E::mul(s_T[aa][bb], d[bb])); }
            }
            w[aa] = acc;
        }
#pragma unroll 4
        for (int64_t q = j0 + lane; q < m; q += 32) {
            T yv = E::ldcg(y + q);
#pragma unroll
            for (int aa = 0; aa < QR_NB; ++aa)
                if (aa < nb) yv = E::sub(yv, E::mul(uget(j0 + aa, q), w[aa]));
            E::stcg(y + q, yv);
        }
    };

    // unblocked factorisation of the nb panel rows [j0, j0 + nb) by the calling CTA
    auto factor_panel = [&](int64_t j0, int nb) {
        if (use_smem) {
            // the nb panel rows are factored in shared memory: one coalesced load, no L2 round trip per step
            __syncthreads();
            for (int64_t e = tid; e < (int64_t)nb * m; e += NT) s_U[e] = E::ldcg(Wt + j0 * m + e);
            __syncthreads();
            for (int aa = 0; aa < nb; ++aa) {
                const int64_t j = j0 + aa;
                T* x = s_U + (int64_t)aa * m;
                double nrm2 = 0.0;
                for (int64_t q = j + lane; q < m; q += 32) nrm2 += E::abs2(x[q]);
                nrm2 = warp_sum(nrm2);
                const T x0 = x[j];
                const double ax0 = sqrt(E::abs2(x0));
                const double nrm = sqrt(nrm2);
                T ph = ax0 > 0.0 ? E::scale(x0, 1.0 / ax0) : E::one();
                const T alpha = E::scale(ph, -nrm);
                const T u0 = E::sub(x0, alpha);
                const double uu = nrm2 - ax0 * ax0 + E::abs2(u0);
                const double tj = (nrm > 0.0 && uu > 0.0) ? 2.0 / uu : 0.0;
                __syncthreads();   // everybody has read row j before it is modified
                if (tid == 0) {
                    tau[j] = tj;
                    uhead[j] = u0;
                    x[j] = tj != 0.0 ? alpha : x0;
                }
                for (int i = aa + 1 + warp; i < nb; i += NW) {
                    T* y = s_U + (int64_t)i * m;
                    T dot = (lane == 0) ? E::mul(E::conj(u0), y[j]) : E::zero();
                    for (int64_t q = j + 1 + lane; q < m; q += 32) dot = E::add(dot, E::mul(E::conj(x[q]), y[q]));
                    dot = E::warp_sum(dot);
                    const T f = E::scale(dot, tj);
                    if (lane == 0) y[j] = E::sub(y[j], E::mul(f, u0));
                    for (int64_t q = j + 1 + lane; q < m; q += 32) y[q] = E::sub(y[q], E::mul(f, x[q]));
                }
                __syncthreads();
            }
            for (int64_t e = tid; e < (int64_t)nb * m; e += NT)
                if (e % m >= j0) E::stcg(Wt + j0 * m + e, s_U[e]);
            __threadfence();
        } else {
            for (int aa = 0; aa < nb; ++aa) {
                const int64_t j = j0 + aa;
                // reflector from x = Wt[j][j:m]; every warp computes the scalars redundantly from the same data
                T* x = Wt + j * m;
                double nrm2 = 0.0;
                for (int64_t q = j + lane; q < m; q += 32) nrm2 += E::abs2(E::ldcg(x + q));
                nrm2 = warp_sum(nrm2);
                const T x0 = E::ldcg(x + j);
                const double ax0 = sqrt(E::abs2(x0));
                const double nrm = sqrt(nrm2);
                T ph = ax0 > 0.0 ? E::scale(x0, 1.0 / ax0) : E::one();
                const T alpha = E::scale(ph, -nrm);       // alpha = -phase(x0) * ||x||
                const T u0 = E::sub(x0, alpha);
                const double uu = nrm2 - ax0 * ax0 + E::abs2(u0);   // u^H u
                const double tj = (nrm > 0.0 && uu > 0.0) ? 2.0 / uu : 0.0;
                __syncthreads();   // everybody has read row j before it is modified
                if (warp == 0 && lane == 0) {
                    tau[j] = tj;
                    uhead[j] = u0;
                    E::stcg(x + j, tj != 0.0 ? alpha : x0);
                }
                // H_j on the remaining panel rows
                for (int64_t i = j + 1 + warp; i < j0 + nb; i += NW) {
                    T* y = Wt + i * m;
                    T dot = (lane == 0) ? E::mul(E::conj(u0), E::ldcg(y + j)) : E::zero();
                    for (int64_t q = j + 1 + lane; q < m; q += 32) dot = E::add(dot, E::mul(E::conj(E::ldcg(x + q)), E::ldcg(y + q)));
                    dot = E::warp_sum(dot);
                    const T f = E::scale(dot, tj);
                    if (lane == 0) E::stcg(y + j, E::sub(E::ldcg(y + j), E::mul(f, u0)));
                    for (int64_t q = j + 1 + lane; q < m; q += 32) E::stcg(y + q, E::sub(E::ldcg(y + q), E::mul(f, E::ldcg(x + q))));
                }
                __threadfence();
                __syncthreads();
            }
        }
    };
    // BLK: outer panel width = what fits shared memory next to nothing else (the staging buffers of the block apply
    // alias it), at most QB_NBO; below 8 rows (m > 3200) the panel stays in global memory (the old inner loop)
    int nbo_max = QB_NBO;
    bool panel_in_smem = false;
    if constexpr (BLK) {
        const int64_t cap = (int64_t)(QB_SMEM_LIMIT / sizeof(double)) / m - QB_NBI;   // + QB_NBI rows of explicit reflectors
        panel_in_smem = cap >= 8;
        if (panel_in_smem && cap < QB_NBO) nbo_max = (int)(cap / QB_NBI) * QB_NBI;
    }
    if constexpr (BLK) {
        double* qb_smem = reinterpret_cast<double*>(qr_smem);
        // factorisation of the outer panel [J0, J0 + nbo) by the calling CTA
        auto factor_outer = [&](int64_t J0, int nbo) {
            const int64_t Jend = J0 + nbo;
            if (panel_in_smem) {
                qr_factor_outer_panel(reinterpret_cast<double*>(Wt), reinterpret_cast<double*>(uhead), tau, J0, nbo, nbo_max, m, qb_smem, prof ? pt : nullptr);
                tick(1);
                return;
            }
            for (int64_t j0 = J0; j0 < Jend; j0 += QR_NB) {
                const int nb = (int)((Jend - j0) < QR_NB ? (Jend - j0) : QR_NB);
                factor_panel(j0, nb);
                __syncthreads();
                tick(1);
                if (j0 + nb < Jend) {           // the rest of the OUTER panel only
                    load_panel(j0, nb);
                    build_T(j0, nb);
                    for (int64_t i = j0 + nb + warp; i < Jend; i += NW) apply_panel(Wt + i * m, j0, nb, true);
                    __threadfence();
                    __syncthreads();
                }
                tick(2);
            }
        };
        // Iteration `it` applies outer panel it-1 to the trailing rows and factors outer panel `it` (one call site each, so
        // that the two big device functions are inlined once and keep their shared-memory address space).
        // Look-ahead: while CTAs 1..G-1 apply panel it-1 to the rows behind panel `it`, CTA 0 applies it to the rows of
        // panel `it` only and factors that panel.  It pays when the others' share of the trailing update (>= two 64-row
        // blocks each) outlasts CTA 0's fixed cost of one small block apply + one panel factorisation (974 x 520: it does
        // not; 3500 x 1800: 115 -> 95 ms).
        const int64_t npanels = (k + nbo_max - 1) / nbo_max;
        const double* Wc = reinterpret_cast<const double*>(Wt);
        const double* uh = reinterpret_cast<const double*>(uhead);
        for (int64_t it = 0; it <= npanels; ++it) {
            const int64_t Jend = it * nbo_max < k ? it * nbo_max : k;                    // first row behind panel it-1
            const int nbo_next = it < npanels ? (int)((k - Jend) < nbo_max ? (k - Jend) : nbo_max) : 0;
            const int64_t Jnext_end = Jend + nbo_next;
            const bool ahead = G > 1 && it >= 1 && nbo_next > 0 && (n - Jnext_end) >= (int64_t)2 * QB_RB * (G - 1);
            if (it >= 1) {
                const int64_t J0 = (it - 1) * nbo_max;
                const int nbo = (int)(Jend - J0);
                int64_t rb = Jend, re = n;
                unsigned c = trank, nc = G;
                if (ahead) {
                    if (trank == 0) { re = Jnext_end; c = 0; nc = 1; }
                    else { rb = Jnext_end; c = trank - 1; nc = G - 1; }
                }
                if (rb < re) qr_block_apply(reinterpret_cast<double*>(Wt), rb, re, Wc, uh, tau, J0, nbo, m, true, qb_smem, c, nc, prof ? pt : nullptr);
                if (nbo_next > 0) {
                    if (ahead) { __threadfence(); __syncthreads(); }
                    else Team<G>::sync();
                }
                tick(3);
            }
            if (nbo_next > 0 && trank == 0) factor_outer(Jend, nbo_next);
            if (nbo_next > 0 || it >= 1) Team<G>::sync();
            tick(2);
        }
    } else {
    for (int64_t j0 = 0; j0 < k; j0 += QR_NB) {
        const int nb = (int)((k - j0) < QR_NB ? (k - j0) : QR_NB);
        if (trank == 0) factor_panel(j0, nb);
        Team<G>::sync();
        if (j0 + nb < n) {
            load_panel(j0, nb);
            build_T(j0, nb);
            for (int64_t i = j0 + nb + gwarp; i < n; i += gwarps) apply_panel(Wt + i * m, j0, nb, true);
        }
        Team<G>::sync();
    }
    }
    // R[j][i] = Wt[i][j] for i >= j, 0 below the diagonal
    for (int64_t e = gtid; e < k * n; e += gthreads) {
        const int64_t j = e / n, i = e % n;
        R[e] = (i >= j) ? E::ldcg(Wt + i * m + j) : E::zero();
    }
    // Qt = [I_k 0] (k x m), then Q <- H_0 ... H_{k-1} [I;0] by backward accumulation
    for (int64_t e = gtid; e < k * m; e += gthreads) E::stcg(&Qt[e], (e / m == e % m) ? E::one() : E::zero());
    Team<G>::sync();
    if constexpr (BLK) {
        double* qb_smem = reinterpret_cast<double*>(qr_smem);
        for (int64_t J0 = ((k - 1) / nbo_max) * nbo_max; J0 >= 0; J0 -= nbo_max) {
            const int nbo = (int)((k - J0) < nbo_max ? (k - J0) : nbo_max);
            // rows c < J0 of Qt are still unit vectors e_c, orthogonal to every u of this panel
            qr_block_apply(reinterpret_cast<double*>(Qt), J0, k, reinterpret_cast<const double*>(Wt),
                           reinterpret_cast<const double*>(uhead), tau, J0, nbo, m, false, qb_smem, trank, G, prof ? pt : nullptr);
            Team<G>::sync();
        }
    } else {
    for (int64_t j0 = ((k - 1) / QR_NB) * QR_NB; j0 >= 0; j0 -= QR_NB) {
        const int nb = (int)((k - j0) < QR_NB ? (k - j0) : QR_NB);
        load_panel(j0, nb);
        build_T(j0, nb);
        // columns c < j0 of Q are still unit vectors e_c, orthogonal to every u of this panel
        for (int64_t cidx = j0 + gwarp; cidx < k; cidx += gwarps) apply_panel(Qt + cidx * m, j0, nb, false);
        Team<G>::sync();
    }
    }
    tick(4);
    for (int64_t e = gtid; e < m * k; e += gthreads) {
        const int64_t q = e / k, cidx = e % k;
        Q[e] = E::ldcg(Qt + cidx * m + q);
    }
    tick(5);
    if (prof && tid == 0 && trank == 0)
        printf("qr team %u m=%lld n=%lld us: load %.1f factor %.1f inner-apply %.1f trailing %.1f form-Q %.1f store %.1f\n",
               Team<G>::id(), (long long)m, (long long)n, pt[0] * 1e-3, pt[1] * 1e-3, pt[2] * 1e-3, pt[3] * 1e-3, pt[4] * 1e-3, pt[5] * 1e-3);
    if (prof && tid == 0 && trank == 0)
        printf("   panel: load %.1f columns %.1f inner-apply %.1f store %.1f | block apply (both phases): pass1 %.1f solve %.1f pass2 %.1f\n",
               pt[6] * 1e-3, pt[7] * 1e-3, pt[8] * 1e-3, pt[9] * 1e-3, pt[10] * 1e-3, pt[11] * 1e-3, pt[12] * 1e-3);
}

template <bool CX>
__global__ void fill_eye_kernel(void* __restrict__ dst_, const int64_t* __restrict__ dims,
                                const int64_t* __restrict__ off, int64_t nmat)
{
    using E = El<CX>;
    using T = typename E::T;
    const int64_t mat = blockIdx.y;
    if (mat >= nmat) return;
    const int64_t m = dims[2 * mat], n = dims[2 * mat + 1];
    T* d = static_cast<T*>(dst_) + off[mat];
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < m * n; e += (int64_t)gridDim.x * blockDim.x)
        d[e] = (e / n == e % n) ? E::one() : E::zero();
}

// ---- host side ----------------------------------------------------------------------------------
enum FzKind { FZ_SVD = 0, FZ_EIGH = 1, FZ_QR = 2 };

template <class K, class... Args>
static cudaError_t launch_team_smem_nt(K kernel, int G, int nteams, size_t smem, int nthreads, cudaStream_t stream, Args... args);

template <class K, class... Args>
static cudaError_t launch_team_smem(K kernel, int G, int nteams, size_t smem, cudaStream_t stream, Args... args)
{
    return launch_team_smem_nt(kernel, G, nteams, smem, FZ_THREADS, stream, args...);
}

template <class K, class... Args>
static cudaError_t launch_team_smem_nt(K kernel, int G, int nteams, size_t smem, int nthreads, cudaStream_t stream, Args... args)
{
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(nteams * G), 1, 1);
    cfg.blockDim = dim3((unsigned)nthreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)G;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

template <class K, class... Args>
static cudaError_t launch_team(K kernel, int G, int nteams, cudaStream_t stream, Args... args)
{
    return launch_team_smem(kernel, G, nteams, 0, stream, args...);
}

constexpr int FZ_BIG_G = 8;          // CTAs per team for large sectors (portable cluster size)
constexpr int64_t FZ_BIG_MIN = 96;   // working rows from which a team is used

// Host driver of the block Jacobi: load phase, sweeps of (sort + 1 + #block-pair rounds) launches, a
// second pass over the noise rows, output phase.
struct BjRounds {
    std::vector<std::vector<BjItem>> rounds;   // rounds[0]: within-block items; rounds[1 + t]: block pairs of round t
    std::vector<BjItem> flat;                  // all rounds back to back
    std::vector<BjWork> gworks, aworks;        // Gram / apply work units of all rounds back to back
    std::vector<size_t> start, gstart, astart; // per round offsets (+ one past the end)
    size_t max_items = 0, max_gworks = 0;
};

// items over the blocks [kb[mat], nb) of every listed matrix
static void bj_build_rounds(BjRounds& R, const std::vector<FzDesc>& descs, const std::vector<int64_t>& mats,
                            const std::vector<int>& rowbase, const std::vector<int>& first_block, int cm, bool with_z)
{
    R.rounds.assign(1, {});
    for (size_t q = 0; q < mats.size(); ++q) {
        const int64_t mat = mats[q];
        const int64_t r = descs[(size_t)mat].r;
        const int nb_all = (int)((r + BJ_B - 1) / BJ_B);
        const int kb = first_block[q];
        const int nb = nb_all - kb;
        if (nb <= 0) continue;
        auto rows_of = [&](int k, int32_t& r0, int32_t& n) { r0 = (kb + k) * BJ_B; n = (int32_t)std::min<int64_t>(BJ_B, r - (int64_t)(kb + k) * BJ_B); };
        for (int k = 0; k < nb; ++k) {
            BjItem it{(int32_t)mat, 0, 0, 0, 0, rowbase[q], 0, 0};
            rows_of(k, it.i0, it.ni);
            if (it.ni > 1) R.rounds[0].push_back(it);
        }
        const int np = nb + (nb & 1);
        if (np < 2) continue;
        if ((int)R.rounds.size() < np) R.rounds.resize((size_t)np);
        for (int t = 0; t < np - 1; ++t)
            for (int k = 0; k < np / 2; ++k) {
                int i = (k == 0) ? np - 1 : (t + k) % (np - 1);
                int j = (k == 0) ? t : (t - k + (np - 1)) % (np - 1);
                if (i > j) std::swap(i, j);
                if (j >= nb) continue;
                BjItem it{(int32_t)mat, 0, 0, 0, 0, rowbase[q], 0, 0};
                rows_of(i, it.i0, it.ni);
                rows_of(j, it.j0, it.nj);
                R.rounds[(size_t)(1 + t)].push_back(it);
            }
    }
    // the most expensive items (longest rows) first within every round; work units of the three kernels
    R.flat.clear(); R.gworks.clear(); R.aworks.clear();
    R.start.clear(); R.gstart.clear(); R.astart.clear();
    R.max_items = R.max_gworks = 0;
    for (auto& rd : R.rounds) {
        std::stable_sort(rd.begin(), rd.end(), [&](const BjItem& x, const BjItem& y) {
            return descs[(size_t)x.mat].c + descs[(size_t)x.mat].r > descs[(size_t)y.mat].c + descs[(size_t)y.mat].r; });
        R.start.push_back(R.flat.size());
        R.gstart.push_back(R.gworks.size());
        R.astart.push_back(R.aworks.size());
        int32_t w = 0;
        for (size_t idx = 0; idx < rd.size(); ++idx) {
            const FzDesc& D = descs[(size_t)rd[idx].mat];
            const int32_t nsl = (int32_t)((D.c * cm + BJ_KS - 1) / BJ_KS);   // cm = 2: real columns of a complex row
            rd[idx].w0 = w; rd[idx].nw = nsl;
            for (int32_t sl = 0; sl < nsl; ++sl) R.gworks.push_back(BjWork{(int32_t)idx, sl});
            w += nsl;
            for (int32_t ch = 0; ch < (int32_t)((D.c * cm + BJ_AC - 1) / BJ_AC); ++ch) R.aworks.push_back(BjWork{(int32_t)idx, ch});
            if (with_z)          // eigh reads its eigenvectors off W: the accumulated left factor is not needed
                for (int32_t ch = 0; ch < (int32_t)((D.r * cm + BJ_AC - 1) / BJ_AC); ++ch) R.aworks.push_back(BjWork{(int32_t)idx, (1 << 24) | ch});
        }
        R.max_items = std::max(R.max_items, rd.size());
        R.max_gworks = std::max(R.max_gworks, (size_t)w);
        R.flat.insert(R.flat.end(), rd.begin(), rd.end());
    }
    R.start.push_back(R.flat.size());
    R.gstart.push_back(R.gworks.size());
    R.astart.push_back(R.aworks.size());
}

template <bool EIGH, bool CX>
static int run_block_jacobi(const void* a, void* o0, double* s_out, void* o2, void* d_ws, double* d_aux,
                            FzDesc* d_desc, const std::vector<FzDesc>& descs, const std::vector<int64_t>& big,
                            const int64_t* sel_big, int64_t* info, cudaStream_t stream)
{
    constexpr int CM = CX ? 2 : 1;
    constexpr size_t SMEM_ROT = CX ? BJ_SMEM_ROT_CX : BJ_SMEM_ROT, SMEM_APPLY = CX ? BJ_SMEM_APPLY_CX : BJ_SMEM_APPLY;
    static PerDeviceOnce configured;
    if (configured.first()) {
        CYB_CHECK_CUDA(cudaFuncSetAttribute(bj_gram_kernel<CX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BJ_SMEM_GRAM));
        CYB_CHECK_CUDA(cudaFuncSetAttribute(bj_rotate_kernel<CX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_ROT));
        CYB_CHECK_CUDA(cudaFuncSetAttribute(bj_apply_kernel<CX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_APPLY));
    }
    const int nbig = (int)big.size();
    const size_t nmat = descs.size();
    cudaError_t err = launch_team(jacobi_svd_kernel<CX, FZ_BIG_G, EIGH, PH_INIT>, FZ_BIG_G, nbig, stream,
                                  a, o0, s_out, o2, d_ws, d_aux, d_desc, sel_big, info);
    if (err != cudaSuccess) { set_error("block jacobi: load launch failed: %s", cudaGetErrorString(err)); return CYB_ERR_CUDA; }
    std::vector<int> rowbase((size_t)nbig + 1, 0);
    for (int q = 0; q < nbig; ++q) rowbase[(size_t)q + 1] = rowbase[(size_t)q] + (int)descs[(size_t)big[(size_t)q]].r;
    const size_t nrows = (size_t)rowbase[(size_t)nbig];
    // Up to four groups of sectors (LPT on the sweep cost r^2 (2c + r)) run their rounds on separate streams: the rotation
    // kernel of a round is latency-bound (one CTA per block pair, tensor pipe idle), so while one group rotates, the Gram /
    // apply kernels of the others keep the tensor pipe busy.  Sectors are independent; only the rounds of ONE sector
    // depend on each other.  chi = 4096 truncated SVD: 71.5 / 66.9 / 65.4 / 64.9 ms with 1 / 2 / 3 / 4 groups
    // (profiles/r2); CYB_BJ_GROUPS overrides.
    constexpr int MAXG = 4;
    int NG = 4;
    if (const char* e = getenv("CYB_BJ_GROUPS")) NG = atoi(e);
    // programmatic dependent launch of the round kernels (bj_pdl_enter), CYB_BJ_PDL=1.  Measured (profiles/r2/svd_groups_r2f.txt):
    // chi = 4096 SVD 58.8 -> 62.1 ms (the early-resident CTAs of the next kernel take shared memory from the other groups'
    // kernels), eigh 22.75 -> 22.19 ms, complex chi = 1024 25.8 -> 25.2 / 13.7 -> 13.1 ms: off by default.
    bool pdl = false;
    if (const char* e = getenv("CYB_BJ_PDL")) pdl = atoi(e) != 0;
    NG = std::max(1, std::min(std::min(NG, MAXG), nbig));
    std::vector<int> group_of((size_t)nbig, 0);
    if (NG >= 2) {
        std::vector<int> idx((size_t)nbig);
        for (int q = 0; q < nbig; ++q) idx[(size_t)q] = q;
        auto cost = [&](int q) { const FzDesc& D = descs[(size_t)big[(size_t)q]]; return (double)D.r * (double)D.r * (2.0 * (double)D.c + (double)D.r); };
        std::stable_sort(idx.begin(), idx.end(), [&](int x, int y) { return cost(x) > cost(y); });
        double load[MAXG] = {0.0, 0.0, 0.0, 0.0};
        for (int q : idx) {
            int g = 0;
            for (int h = 1; h < NG; ++h) if (load[h] < load[g]) g = h;
            group_of[(size_t)q] = g; load[g] += cost(q);
        }
    }
    std::vector<int64_t> gmats[MAXG];
    std::vector<int> growbase[MAXG];
    for (int q = 0; q < nbig; ++q) {
        const int g = group_of[(size_t)q];
        gmats[g].push_back(big[(size_t)q]);
        growbase[g].push_back(rowbase[(size_t)q]);
    }
    BjRounds R1[MAXG], R2[MAXG];
    size_t max_flat = 1, max_gw = 1, max_aw = 1, max_items = 1, max_gworks = 1;
    for (int g = 0; g < NG; ++g) {
        bj_build_rounds(R1[g], descs, gmats[g], growbase[g], std::vector<int>(gmats[g].size(), 0), CM, !EIGH);
        max_flat = std::max(max_flat, R1[g].flat.size()); max_gw = std::max(max_gw, R1[g].gworks.size());
        max_aw = std::max(max_aw, R1[g].aworks.size()); max_items = std::max(max_items, R1[g].max_items);
        max_gworks = std::max(max_gworks, R1[g].max_gworks);
    }
    // device scratch, per group: items | Gram works | apply works (pass-2 lists are never longer than pass-1 ones) | skip |
    //                 Gpart | Q;   shared: active | flags | nnoise | rowbase | rowmap | norm2 | floor2
    const size_t items_bytes = align_up(sizeof(BjItem) * max_flat, 256);
    const size_t gw_bytes = align_up(sizeof(BjWork) * max_gw, 256);
    const size_t aw_bytes = align_up(sizeof(BjWork) * max_aw, 256);
    const size_t skip_bytes = align_up(sizeof(int) * max_items, 256);
    const size_t gpart_bytes = align_up(sizeof(double) * BJ_R * BJ_R * CM * max_gworks, 256);
    const size_t q_bytes = align_up(sizeof(double) * BJ_R * BJ_R * CM * max_items, 256);
    const size_t group_bytes = items_bytes + gw_bytes + aw_bytes + skip_bytes + gpart_bytes + q_bytes;
    const size_t ints_bytes = align_up(sizeof(int) * nmat, 256);
    const size_t rb_bytes = align_up(sizeof(int) * ((size_t)nbig + 1), 256);
    const size_t map_bytes = align_up(sizeof(int) * std::max<size_t>(nrows, 1), 256);
    const size_t nrm_bytes = align_up(sizeof(double) * std::max<size_t>(nrows, 1), 256);
    const size_t fl_bytes = align_up(sizeof(double) * nmat, 256);
    Scratch scr(NG * group_bytes + 3 * ints_bytes + rb_bytes + map_bytes + nrm_bytes + fl_bytes, stream);
    if (!scr.ptr) return CYB_ERR_CUDA;
    size_t o = 0;
    BjItem* d_items[MAXG]; BjWork* d_gworks[MAXG]; BjWork* d_aworks[MAXG]; int* d_skip[MAXG]; double* d_gpart[MAXG]; double* d_qbuf[MAXG];
    for (int g = 0; g < NG; ++g) {
        d_items[g] = scr.as<BjItem>(o); o += items_bytes;
        d_gworks[g] = scr.as<BjWork>(o); o += gw_bytes;
        d_aworks[g] = scr.as<BjWork>(o); o += aw_bytes;
        d_skip[g] = scr.as<int>(o); o += skip_bytes;
        d_gpart[g] = scr.as<double>(o); o += gpart_bytes;
        d_qbuf[g] = scr.as<double>(o); o += q_bytes;
    }
    int* d_active = scr.as<int>(o); o += ints_bytes;
    int* d_flags = scr.as<int>(o); o += ints_bytes;
    int* d_nnoise = scr.as<int>(o); o += ints_bytes;
    int* d_rowbase = scr.as<int>(o); o += rb_bytes;
    int* d_rowmap = scr.as<int>(o); o += map_bytes;
    double* d_norm2 = scr.as<double>(o); o += nrm_bytes;
    double* d_floor2 = scr.as<double>(o);
    std::vector<int> active(nmat, 0), flags(nmat, 0), nnoise(nmat, 0);
    std::vector<int64_t> sweeps(nmat, 0);
    for (int64_t mat : big) active[(size_t)mat] = 1;
    CYB_CHECK_CUDA(cudaMemcpyAsync(d_rowbase, rowbase.data(), sizeof(int) * rowbase.size(), cudaMemcpyHostToDevice, stream));
    CYB_CHECK_CUDA(cudaMemsetAsync(d_floor2, 0, sizeof(double) * nmat, stream));
    // the second stream and its fork / join events (per device, created once)
    static cudaStream_t side_stream[64][MAXG] = {};
    static cudaEvent_t ev_fork[64] = {}, ev_join[64][MAXG] = {};
    int devid = 0;
    CYB_CHECK_CUDA(cudaGetDevice(&devid));
    devid &= 63;
    if (NG >= 2 && !ev_fork[devid]) {
        CYB_CHECK_CUDA(cudaEventCreateWithFlags(&ev_fork[devid], cudaEventDisableTiming));
        for (int g = 1; g < MAXG; ++g) {
            CYB_CHECK_CUDA(cudaStreamCreateWithFlags(&side_stream[devid][g], cudaStreamNonBlocking));
            CYB_CHECK_CUDA(cudaEventCreateWithFlags(&ev_join[devid][g], cudaEventDisableTiming));
        }
    }
    cudaStream_t gstream[MAXG] = {stream, stream, stream, stream};
    for (int g = 1; g < NG; ++g) gstream[g] = side_stream[devid][g];
    const int norm_grid = sm_count() * 4;
    auto sort_rows = [&](int init) {
        bj_norms_kernel<<<norm_grid, 256, 0, stream>>>(static_cast<const double*>(d_ws), d_desc, sel_big, d_rowbase, nbig, d_norm2, CM);
        bj_sort_kernel<<<nbig, 1024, 0, stream>>>(d_desc, sel_big, d_rowbase, d_norm2, d_rowmap, d_floor2, d_nnoise, d_active, init);
    };
    // runs sweeps over the item lists of R[0..NG) until no item rotates
    auto run_sweeps = [&](const BjRounds* R, bool phase1) -> int {
        bool have = false;
        size_t nrounds = 0;
        for (int g = 0; g < NG; ++g) {
            if (R[g].flat.empty()) continue;
            have = true;
            nrounds = std::max(nrounds, R[g].rounds.size());
            CYB_CHECK_CUDA(cudaMemcpyAsync(d_items[g], R[g].flat.data(), sizeof(BjItem) * R[g].flat.size(), cudaMemcpyHostToDevice, stream));
            CYB_CHECK_CUDA(cudaMemcpyAsync(d_gworks[g], R[g].gworks.data(), sizeof(BjWork) * R[g].gworks.size(), cudaMemcpyHostToDevice, stream));
            CYB_CHECK_CUDA(cudaMemcpyAsync(d_aworks[g], R[g].aworks.data(), sizeof(BjWork) * R[g].aworks.size(), cudaMemcpyHostToDevice, stream));
        }
        if (!have) return CYB_OK;
        bool any = false;
        for (size_t m = 0; m < nmat; ++m) any = any || active[m];
        for (int sweep = 0; sweep < FZ_MAX_SWEEPS && any; ++sweep) {
            CYB_CHECK_CUDA(cudaMemcpyAsync(d_active, active.data(), sizeof(int) * nmat, cudaMemcpyHostToDevice, stream));
            CYB_CHECK_CUDA(cudaMemsetAsync(d_flags, 0, sizeof(int) * nmat, stream));
            if (phase1) sort_rows(sweep == 0);
            if (NG >= 2) {
                CYB_CHECK_CUDA(cudaEventRecord(ev_fork[devid], stream));
                for (int g = 1; g < NG; ++g) CYB_CHECK_CUDA(cudaStreamWaitEvent(gstream[g], ev_fork[devid], 0));
            }
            for (size_t rd = 0; rd < nrounds; ++rd) {
                for (int g = 0; g < NG; ++g) {
                    if (rd >= R[g].rounds.size()) continue;
                    bool live = false;   // skip rounds whose items all belong to converged matrices
                    for (const BjItem& it : R[g].rounds[rd]) if (active[(size_t)it.mat]) { live = true; break; }
                    if (!live) continue;
                    const unsigned ni = (unsigned)R[g].rounds[rd].size();
                    const unsigned ng = (unsigned)(R[g].gstart[rd + 1] - R[g].gstart[rd]), na = (unsigned)(R[g].astart[rd + 1] - R[g].astart[rd]);
                    cudaError_t le = bj_launch(pdl, bj_gram_kernel<CX>, ng, BJ_THREADS, BJ_SMEM_GRAM, gstream[g],
                        static_cast<const double*>(d_ws), (const FzDesc*)d_desc, (const BjItem*)(d_items[g] + R[g].start[rd]),
                        (const BjWork*)(d_gworks[g] + R[g].gstart[rd]), (const int*)d_active, (const int*)d_rowmap, d_gpart[g]);
                    if (le == cudaSuccess)
                        le = bj_launch(pdl, bj_rotate_kernel<CX>, ni, BJ_RT, SMEM_ROT, gstream[g],
                            (const FzDesc*)d_desc, (const BjItem*)(d_items[g] + R[g].start[rd]), (const int*)d_active, d_flags,
                            (const double*)d_floor2, phase1 ? 1 : 0, (const double*)d_gpart[g], d_qbuf[g], d_skip[g]);
                    if (le == cudaSuccess)
                        le = bj_launch(pdl, bj_apply_kernel<CX>, na, BJ_THREADS, SMEM_APPLY, gstream[g],
                            static_cast<double*>(d_ws), (const FzDesc*)d_desc, (const BjItem*)(d_items[g] + R[g].start[rd]),
                            (const BjWork*)(d_aworks[g] + R[g].astart[rd]), (const int*)d_rowmap, (const double*)d_qbuf[g], (const int*)d_skip[g]);
                    if (le != cudaSuccess) { set_error("block jacobi: round launch failed: %s", cudaGetErrorString(le)); return CYB_ERR_CUDA; }
                }
            }
            CYB_LAUNCH_CHECK();
            for (int g = 1; g < NG; ++g) {
                CYB_CHECK_CUDA(cudaEventRecord(ev_join[devid][g], gstream[g]));
                CYB_CHECK_CUDA(cudaStreamWaitEvent(stream, ev_join[devid][g], 0));
            }
            CYB_CHECK_CUDA(cudaMemcpyAsync(flags.data(), d_flags, sizeof(int) * nmat, cudaMemcpyDeviceToHost, stream));
            CYB_CHECK_CUDA(cudaStreamSynchronize(stream));
            any = false;
            for (size_t m = 0; m < nmat; ++m) {
                if (!active[m]) continue;
                sweeps[m] += 1;
                active[m] = flags[m];
                any = any || flags[m];
            }
        }
        return CYB_OK;
    };
    // ---- pass 1: all rows, pairs of two noise rows left alone ----
    int rc = run_sweeps(R1, true);
    if (rc != CYB_OK) return rc;
    bool failed = false;
    for (size_t m = 0; m < nmat; ++m) failed = failed || active[m];
    const bool debug = getenv("CYB_BJ_DEBUG") != nullptr;          // sweeps per sector after each pass, on stderr
    if (debug)
        for (int64_t mat : big)
            fprintf(stderr, "block jacobi: sector %lld (%lld x %lld) pass 1: %lld sweeps\n", (long long)mat,
                    (long long)descs[(size_t)mat].r, (long long)descs[(size_t)mat].c, (long long)sweeps[(size_t)mat]);
    // ---- pass 2: orthogonalise the noise rows among themselves (trailing blocks of the sorted order) ----
    if (!failed) {
        for (int64_t mat : big) active[(size_t)mat] = 1;
        CYB_CHECK_CUDA(cudaMemcpyAsync(d_active, active.data(), sizeof(int) * nmat, cudaMemcpyHostToDevice, stream));
        sort_rows(0);
        CYB_LAUNCH_CHECK();
        CYB_CHECK_CUDA(cudaMemcpyAsync(nnoise.data(), d_nnoise, sizeof(int) * nmat, cudaMemcpyDeviceToHost, stream));
        CYB_CHECK_CUDA(cudaStreamSynchronize(stream));
        for (int g = 0; g < NG; ++g) {
            std::vector<int> first_block(gmats[g].size(), 0);
            for (size_t q = 0; q < gmats[g].size(); ++q) {
                const int64_t mat = gmats[g][q];
                const int64_t r = descs[(size_t)mat].r;
                const int nn = nnoise[(size_t)mat];
                active[(size_t)mat] = nn > 1;
                first_block[q] = nn > 1 ? (int)((r - nn) / BJ_B) : (int)((r + BJ_B - 1) / BJ_B);
            }
            bj_build_rounds(R2[g], descs, gmats[g], growbase[g], first_block, CM, !EIGH);
        }
        rc = run_sweeps(R2, false);
        if (rc != CYB_OK) return rc;
        if (debug)
            for (int64_t mat : big)
                fprintf(stderr, "block jacobi: sector %lld pass 1 + 2: %lld sweeps, %d rows at the noise floor\n", (long long)mat,
                        (long long)sweeps[(size_t)mat], nnoise[(size_t)mat]);
    }
    err = launch_team(jacobi_svd_kernel<CX, FZ_BIG_G, EIGH, PH_FINISH>, FZ_BIG_G, nbig, stream,
                      a, o0, s_out, o2, d_ws, d_aux, d_desc, sel_big, info);
    if (err != cudaSuccess) { set_error("block jacobi: output launch failed: %s", cudaGetErrorString(err)); return CYB_ERR_CUDA; }
    // info: sweeps used (negative: not converged within FZ_MAX_SWEEPS)
    std::vector<int64_t> vals((size_t)nbig);
    for (int q = 0; q < nbig; ++q) {
        const size_t mat = (size_t)big[(size_t)q];
        vals[(size_t)q] = std::max<int64_t>(sweeps[mat], 1) * (active[mat] ? -1 : 1);
        CYB_CHECK_CUDA(cudaMemcpyAsync(info + mat, &vals[(size_t)q], sizeof(int64_t), cudaMemcpyHostToDevice, stream));
    }
    CYB_CHECK_CUDA(cudaStreamSynchronize(stream));   // the host vectors must outlive the copies
    return CYB_OK;
}

static int run_factorize(FzKind kind, const void* a, void* o0, double* o1d, void* o1, void* o2,
                         const int64_t* desc_host, int64_t nmat, int dtype, int64_t* info,
                         cudaStream_t stream)
{
    if (nmat == 0) return CYB_OK;
    const bool cx = dtype == CYB_C128;
    const size_t esz = cx ? 16 : 8;
    std::vector<FzDesc> descs((size_t)nmat);
    std::vector<int64_t> small, big;
    int64_t ws_elems = 0, aux_doubles = 0;
    for (int64_t i = 0; i < nmat; ++i) {
        const int64_t* h = desc_host + 6 * i;
        FzDesc& D = descs[(size_t)i];
        D.m = h[0]; D.n = h[1]; D.a_off = h[2]; D.o0_off = h[3]; D.o1_off = h[4]; D.o2_off = h[5];
        D.shift = 0.0;
        if (D.m < 0 || D.n < 0) { set_error("factorize: negative dims"); return CYB_ERR_INVALID; }
        if (kind == FZ_EIGH && D.m != D.n) { set_error("eigh: block %lld is not square", (long long)i); return CYB_ERR_INVALID; }
        if (D.m == 0 || D.n == 0) { D.r = D.c = 0; D.w_off = D.z_off = D.aux_off = 0; D.transposed = 0; continue; }
        if (kind == FZ_QR) {
            const int64_t k = D.m < D.n ? D.m : D.n;
            D.r = D.n; D.c = D.m; D.transposed = 1;
            D.w_off = ws_elems; ws_elems += align_up((size_t)(D.n * D.m), 16);
            D.z_off = ws_elems; ws_elems += align_up((size_t)(k * D.m), 16);
            D.aux_off = aux_doubles; aux_doubles += (int64_t)align_up((size_t)(3 * k + 8), 16);
            (k >= FZ_BIG_MIN ? big : small).push_back(i);
        } else {
            D.transposed = (D.m > D.n) ? 1 : 0;
            D.r = D.transposed ? D.n : D.m;
            D.c = D.transposed ? D.m : D.n;
            D.w_off = ws_elems; ws_elems += align_up((size_t)(D.r * D.c), 16);
            D.z_off = ws_elems; ws_elems += align_up((size_t)(D.r * D.r), 16);
            D.aux_off = aux_doubles; aux_doubles += (int64_t)align_up((size_t)svd_aux_doubles(D.r, D.c), 16);
            (D.r >= FZ_BIG_MIN ? big : small).push_back(i);
        }
    }
    const size_t desc_bytes = align_up(sizeof(FzDesc) * (size_t)nmat, 256);
    const size_t sel_bytes = align_up(sizeof(int64_t) * (size_t)(nmat + 1), 256);
    const size_t ws_bytes = align_up((size_t)ws_elems * esz + 16, 256);
    const size_t aux_bytes = align_up((size_t)aux_doubles * 8 + 16, 256);
    Scratch scr(desc_bytes + sel_bytes + ws_bytes + aux_bytes, stream);
    if (!scr.ptr) return CYB_ERR_CUDA;
    FzDesc* d_desc = scr.as<FzDesc>();
    int64_t* d_sel = scr.as<int64_t>(desc_bytes);
    void* d_ws = scr.as<char>(desc_bytes + sel_bytes);
    double* d_aux = scr.as<double>(desc_bytes + sel_bytes + ws_bytes);
    std::vector<int64_t> sel(small);
    sel.insert(sel.end(), big.begin(), big.end());
    CYB_CHECK_CUDA(cudaMemcpyAsync(d_desc, descs.data(), sizeof(FzDesc) * (size_t)nmat, cudaMemcpyHostToDevice, stream));
    if (!sel.empty())
        CYB_CHECK_CUDA(cudaMemcpyAsync(d_sel, sel.data(), sizeof(int64_t) * sel.size(), cudaMemcpyHostToDevice, stream));
    // the host vectors go out of scope at return: pageable cudaMemcpyAsync has already staged them
    const int64_t* sel_small = d_sel;
    const int64_t* sel_big = d_sel + small.size();
    cudaError_t err = cudaSuccess;
#define FZ_LAUNCH(KERN, G, NT, ...) \
    do { if ((NT) > 0) { err = launch_team(KERN, G, (int)(NT), stream, __VA_ARGS__); if (err != cudaSuccess) { set_error("factorize launch failed: %s", cudaGetErrorString(err)); return CYB_ERR_CUDA; } } } while (0)
#define FZ_LAUNCH_SMEM(KERN, G, NT, SMEM, ...) \
    do { if ((NT) > 0) { err = launch_team_smem(KERN, G, (int)(NT), SMEM, stream, __VA_ARGS__); if (err != cudaSuccess) { set_error("factorize launch failed: %s", cudaGetErrorString(err)); return CYB_ERR_CUDA; } } } while (0)
    // QR: the QR_NB x m panel lives in shared memory when it fits (else the kernels work through L2)
    size_t qr_smem_small = 0, qr_smem_big = 0;
    bool qr_panel_big = false;
    const int qr_prof = (kind == FZ_QR && getenv("CYB_QR_PROFILE")) ? 2 : 0;
    if (kind == FZ_QR) {
        int64_t m_small = 0, m_big = 0;
        for (int64_t i : small) m_small = std::max(m_small, descs[(size_t)i].m);
        for (int64_t i : big) m_big = std::max(m_big, descs[(size_t)i].m);
        const size_t lim = 200 * 1024;
        qr_smem_small = (size_t)QR_NB * (size_t)m_small * esz;
        qr_smem_big = (size_t)QR_NB * (size_t)m_big * esz;
        if (qr_smem_small > lim) qr_smem_small = 0;
        if (qr_smem_big > lim) qr_smem_big = 0;
        qr_panel_big = qr_smem_big > 0;
        // real large sectors: the outer panel (factored in shared memory) and the operand staging area of the DMMA block
        // apply share one region of QB_SMEM_LIMIT bytes; the QR_NB x m inner panel of the global-memory fallback too
        if (!cx) {
            static_assert(QB_SMEM_DOUBLES * sizeof(double) <= QB_SMEM_LIMIT, "block apply staging exceeds the shared-memory budget");
            qr_panel_big = (size_t)QR_NB * (size_t)m_big * esz <= QB_SMEM_LIMIT;
            qr_smem_big = QB_SMEM_LIMIT;
        }
    }
    if (kind != FZ_QR && !big.empty()) {
        // large sectors: block Jacobi (block_jacobi.cuh) between the load and the output phases
        const int rc = (kind == FZ_EIGH)
            ? (cx ? run_block_jacobi<true, true>(a, o0, o1d, o2, d_ws, d_aux, d_desc, descs, big, sel_big, info, stream)
                  : run_block_jacobi<true, false>(a, o0, o1d, o2, d_ws, d_aux, d_desc, descs, big, sel_big, info, stream))
            : (cx ? run_block_jacobi<false, true>(a, o0, o1d, o2, d_ws, d_aux, d_desc, descs, big, sel_big, info, stream)
                  : run_block_jacobi<false, false>(a, o0, o1d, o2, d_ws, d_aux, d_desc, descs, big, sel_big, info, stream));
        if (rc != CYB_OK) return rc;
        big.clear();
    }
    if (kind == FZ_SVD) {
        if (cx) {
            FZ_LAUNCH((jacobi_svd_kernel<true, 1, false, PH_ALL>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<true, FZ_BIG_G, false, PH_ALL>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        } else {
            FZ_LAUNCH((jacobi_svd_kernel<false, 1, false, PH_ALL>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<false, FZ_BIG_G, false, PH_ALL>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        }
    } else if (kind == FZ_EIGH) {
        if (cx) {
            FZ_LAUNCH((jacobi_svd_kernel<true, 1, true, PH_ALL>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<true, FZ_BIG_G, true, PH_ALL>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        } else {
            FZ_LAUNCH((jacobi_svd_kernel<false, 1, true, PH_ALL>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<false, FZ_BIG_G, true, PH_ALL>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        }
    } else {
        if (cx) {
            FZ_LAUNCH_SMEM((householder_qr_kernel<true, 1, false>), 1, small.size(), qr_smem_small, a, o0, o1, d_ws, d_aux, d_desc, sel_small, (int)(qr_smem_small > 0));
            FZ_LAUNCH_SMEM((householder_qr_kernel<true, FZ_BIG_G, false>), FZ_BIG_G, big.size(), qr_smem_big, a, o0, o1, d_ws, d_aux, d_desc, sel_big, (int)qr_panel_big | qr_prof);
        } else {
            FZ_LAUNCH_SMEM((householder_qr_kernel<false, 1, false>), 1, small.size(), qr_smem_small, a, o0, o1, d_ws, d_aux, d_desc, sel_small, (int)(qr_smem_small > 0));
            if (!big.empty()) {
                err = launch_team_smem_nt(householder_qr_kernel<false, FZ_BIG_G, true>, FZ_BIG_G, (int)big.size(), qr_smem_big, QB_THREADS,
                                          stream, a, o0, o1, d_ws, d_aux, d_desc, sel_big, (int)qr_panel_big | qr_prof);
                if (err != cudaSuccess) { set_error("factorize launch failed: %s", cudaGetErrorString(err)); return CYB_ERR_CUDA; }
            }
        }
    }
#undef FZ_LAUNCH
#undef FZ_LAUNCH_SMEM
    return CYB_OK;
}

}  // namespace cyb

using namespace cyb;

extern "C" {

int cyb_batched_svd(const void* a, void* u, double* s, void* vh, const int64_t* desc_host,
                    int64_t nmat, int dtype, int64_t* info, cyb_stream_t stream)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "batched_svd: bad dtype %d", dtype);
    CYB_REQUIRE(nmat >= 0, "batched_svd: nmat < 0");
    return run_factorize(FZ_SVD, a, u, s, nullptr, vh, desc_host, nmat, dtype, info, (cudaStream_t)stream);
}

int cyb_batched_eigh(const void* a, double* w, void* v, const int64_t* desc_host, int64_t nmat,
                     int dtype, int64_t* info, cyb_stream_t stream)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "batched_eigh: bad dtype %d", dtype);
    CYB_REQUIRE(nmat >= 0, "batched_eigh: nmat < 0");
    return run_factorize(FZ_EIGH, a, nullptr, w, nullptr, v, desc_host, nmat, dtype, info, (cudaStream_t)stream);
}

int cyb_batched_qr(const void* a, void* q, void* r, const int64_t* desc_host, int64_t nmat,
                   int dtype, cyb_stream_t stream)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "batched_qr: bad dtype %d", dtype);
    CYB_REQUIRE(nmat >= 0, "batched_qr: nmat < 0");
    return run_factorize(FZ_QR, a, q, nullptr, r, nullptr, desc_host, nmat, dtype, nullptr, (cudaStream_t)stream);
}

int cyb_fill_eye(void* dst, const int64_t* dims, const int64_t* off, int64_t nmat, int dtype,
                 cyb_stream_t stream)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "fill_eye: bad dtype %d", dtype);
    if (nmat <= 0) return CYB_OK;
    CYB_REQUIRE(nmat <= 65535, "fill_eye: too many blocks (%lld)", (long long)nmat);
    dim3 grid(64, (unsigned)nmat);
    if (dtype == CYB_C128) fill_eye_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(dst, dims, off, nmat);
    else fill_eye_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(dst, dims, off, nmat);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

}  // extern "C"
