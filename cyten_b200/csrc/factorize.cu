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
//           left/right singular vectors coincide with the eigenvectors; lambda = S - sigma.
//           Replaces block_backend->eigh (abelian.cpp:1781; numpy.cpp:658-680 np.linalg.eigh).
//  * QR   : Householder QR on the transposed copy (columns of A become contiguous rows), one warp
//           per trailing column, then backward accumulation of Q.
//           Replaces block_backend->matrix_qr(block, full=false) (abelian.cpp:3123;
//           numpy.cpp:1236-1245 scipy.linalg.qr economic).
//  Gauge: singular/eigen vectors are defined up to phases (and rotations in degenerate spaces), R's
//  diagonal up to phases; parity with the reference is on S / W values, reconstruction and isometry.
#include "common.cuh"
#include <vector>
#include <cmath>

namespace cyb {

constexpr int FZ_THREADS = 1024;
constexpr int FZ_WARPS = FZ_THREADS / 32;
constexpr int FZ_MAX_SWEEPS = 60;
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

template <bool CX, int G, bool EIGH>
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

    // ---- load working copy, Z = I ----
    if (EIGH) {
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
    for (int64_t e = gtid; e < r * r; e += gthreads) E::stcg(&Z[e], (e / r == e % r) ? E::one() : E::zero());
    for (int64_t e = gtid; e < FZ_MAX_SWEEPS + 1; e += gthreads) rot_count[e] = 0ull;
    Team<G>::sync();

    // ---- Jacobi sweeps ----
    const double tol = FZ_EPS * sqrt((double)c);
    const int64_t rp = r + (r & 1);          // players (padded to even)
    const int64_t rounds = rp - 1, npairs = rp / 2;
    int sweeps_done = 0;
    bool converged = (r < 2);
    for (int sweep = 0; sweep < FZ_MAX_SWEEPS && !converged; ++sweep) {
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
                T* zx = Z + i * r;
                T* zy = Z + j * r;
                for (int64_t q = lane; q < r; q += 32) {
                    const T xv = E::ldcg(zx + q), yv = E::mul(phase, E::ldcg(zy + q));
                    E::stcg(zx + q, E::sub(E::scale(xv, cs), E::scale(yv, sn)));
                    E::stcg(zy + q, E::add(E::scale(xv, sn), E::scale(yv, cs)));
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
        // w (ascending) and V[j][p] = conj(Z[i][j]) with p = pos[i]
        T* V = static_cast<T*>(o2_) + D.o2_off;
        for (int64_t i = gtid; i < r; i += gthreads) s_out[D.o1_off + __ldcg(&pos[i])] = __ldcg(&S_raw[i]) - sigma;
        for (int64_t e = gtid; e < r * r; e += gthreads) {
            const int64_t j = e / r, i = e % r;   // consecutive threads -> consecutive i (scattered columns)
            V[j * r + __ldcg(&pos[i])] = E::conj(E::ldcg(Z + i * r + j));
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
    if (gtid == 0) info[mat] = converged ? (sweeps_done > 0 ? sweeps_done : 1) : -(sweeps_done > 0 ? sweeps_done : 1);
    (void)m;
}

// ---- QR ---------------------------------------------------------------------------------------
// Working copy Wt = A^T (n x m): row i of Wt is column i of A.  Hermitian Householder reflectors
// H_j = I - tau_j u_j u_j^H (tau real), u_j stored in Wt[j][j+1:m] with its head in uhead[j].
template <bool CX, int G>
__global__ void __launch_bounds__(FZ_THREADS, 1)
householder_qr_kernel(const void* __restrict__ a_, void* __restrict__ q_, void* __restrict__ r_,
                      void* __restrict__ ws_, double* __restrict__ aux_, FzDesc* __restrict__ descs,
                      const int64_t* __restrict__ sel)
{
    using E = El<CX>;
    using T = typename E::T;
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
    const unsigned trank = Team<G>::rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t gthreads = (int64_t)G * FZ_THREADS, gtid = (int64_t)trank * FZ_THREADS + tid;
    const int64_t gwarps = (int64_t)G * FZ_WARPS, gwarp = (int64_t)trank * FZ_WARPS + warp;

    for (int64_t e = gtid; e < n * m; e += gthreads) {
        const int64_t i = e / m, q = e % m;
        E::stcg(&Wt[e], A[q * n + i]);
    }
    Team<G>::sync();

    for (int64_t j = 0; j < k; ++j) {
        // reflector from x = Wt[j][j:m]  (one warp of the team's CTA 0... every CTA's warp 0
        // computes it redundantly from the same data so no extra barrier is needed)
        T* x = Wt + j * m;
        double nrm2 = 0.0;
        for (int64_t q = j + lane; q < m; q += 32) nrm2 += E::abs2(E::ldcg(x + q));
        nrm2 = warp_sum(nrm2);
        const T x0 = E::ldcg(x + j);
        const double ax0 = sqrt(E::abs2(x0));
        const double nrm = sqrt(nrm2);
        // alpha = -phase(x0) * ||x||
        T ph = ax0 > 0.0 ? E::scale(x0, 1.0 / ax0) : E::one();
        const T alpha = E::scale(ph, -nrm);
        const T u0 = E::sub(x0, alpha);
        // u^H u = ||x||^2 - |x0|^2 + |u0|^2
        const double uu = nrm2 - ax0 * ax0 + E::abs2(u0);
        const double tj = (nrm > 0.0 && uu > 0.0) ? 2.0 / uu : 0.0;
        Team<G>::sync();   // everybody has read row j before it is modified
        if (gwarp == 0 && lane == 0) {
            tau[j] = tj;
            uhead[j] = u0;
            E::stcg(x + j, tj != 0.0 ? alpha : x0);
        }
        // apply H_j to the trailing columns (rows i > j of Wt): y <- y - tau u (u^H y)
        for (int64_t i = j + 1 + gwarp; i < n; i += gwarps) {
            T* y = Wt + i * m;
            T dot = (lane == 0) ? E::mul(E::conj(u0), E::ldcg(y + j)) : E::zero();
            for (int64_t q = j + 1 + lane; q < m; q += 32) dot = E::add(dot, E::mul(E::conj(E::ldcg(x + q)), E::ldcg(y + q)));
            dot = E::warp_sum(dot);
            const T f = E::scale(dot, tj);
            if (lane == 0) E::stcg(y + j, E::sub(E::ldcg(y + j), E::mul(f, u0)));
            for (int64_t q = j + 1 + lane; q < m; q += 32) E::stcg(y + q, E::sub(E::ldcg(y + q), E::mul(f, E::ldcg(x + q))));
        }
        Team<G>::sync();
    }
    // R[j][i] = Wt[i][j] for i >= j, 0 below the diagonal
    for (int64_t e = gtid; e < k * n; e += gthreads) {
        const int64_t j = e / n, i = e % n;
        R[e] = (i >= j) ? E::ldcg(Wt + i * m + j) : E::zero();
    }
    // Qt = [I_k 0] (k x m), then Q <- H_0 ... H_{k-1} [I;0] by backward accumulation
    for (int64_t e = gtid; e < k * m; e += gthreads) E::stcg(&Qt[e], (e / m == e % m) ? E::one() : E::zero());
    Team<G>::sync();
    for (int64_t j = k - 1; j >= 0; --j) {
        const double tj = __ldcg(&tau[j]);
        const T u0 = E::ldcg(&uhead[j]);
        const T* x = Wt + j * m;
        if (tj != 0.0) {
            for (int64_t cidx = j + gwarp; cidx < k; cidx += gwarps) {
                T* y = Qt + cidx * m;
                T dot = (lane == 0) ? E::mul(E::conj(u0), E::ldcg(y + j)) : E::zero();
                for (int64_t q = j + 1 + lane; q < m; q += 32) dot = E::add(dot, E::mul(E::conj(E::ldcg(x + q)), E::ldcg(y + q)));
                dot = E::warp_sum(dot);
                const T f = E::scale(dot, tj);
                if (lane == 0) E::stcg(y + j, E::sub(E::ldcg(y + j), E::mul(f, u0)));
                for (int64_t q = j + 1 + lane; q < m; q += 32) E::stcg(y + q, E::sub(E::ldcg(y + q), E::mul(f, E::ldcg(x + q))));
            }
        }
        Team<G>::sync();
    }
    for (int64_t e = gtid; e < m * k; e += gthreads) {
        const int64_t q = e / k, cidx = e % k;
        Q[e] = E::ldcg(Qt + cidx * m + q);
    }
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
static cudaError_t launch_team(K kernel, int G, int nteams, cudaStream_t stream, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(nteams * G), 1, 1);
    cfg.blockDim = dim3(FZ_THREADS, 1, 1);
    cfg.dynamicSmemBytes = 0;
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

constexpr int FZ_BIG_G = 8;          // CTAs per team for large sectors (portable cluster size)
constexpr int64_t FZ_BIG_MIN = 96;   // working rows from which a team is used

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
    if (kind == FZ_SVD) {
        if (cx) {
            FZ_LAUNCH((jacobi_svd_kernel<true, 1, false>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<true, FZ_BIG_G, false>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        } else {
            FZ_LAUNCH((jacobi_svd_kernel<false, 1, false>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<false, FZ_BIG_G, false>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        }
    } else if (kind == FZ_EIGH) {
        if (cx) {
            FZ_LAUNCH((jacobi_svd_kernel<true, 1, true>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<true, FZ_BIG_G, true>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        } else {
            FZ_LAUNCH((jacobi_svd_kernel<false, 1, true>), 1, small.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_small, info);
            FZ_LAUNCH((jacobi_svd_kernel<false, FZ_BIG_G, true>), FZ_BIG_G, big.size(), a, o0, o1d, o2, d_ws, d_aux, d_desc, sel_big, info);
        }
    } else {
        if (cx) {
            FZ_LAUNCH((householder_qr_kernel<true, 1>), 1, small.size(), a, o0, o1, d_ws, d_aux, d_desc, sel_small);
            FZ_LAUNCH((householder_qr_kernel<true, FZ_BIG_G>), FZ_BIG_G, big.size(), a, o0, o1, d_ws, d_aux, d_desc, sel_big);
        } else {
            FZ_LAUNCH((householder_qr_kernel<false, 1>), 1, small.size(), a, o0, o1, d_ws, d_aux, d_desc, sel_small);
            FZ_LAUNCH((householder_qr_kernel<false, FZ_BIG_G>), FZ_BIG_G, big.size(), a, o0, o1, d_ws, d_aux, d_desc, sel_big);
        }
    }
#undef FZ_LAUNCH
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
