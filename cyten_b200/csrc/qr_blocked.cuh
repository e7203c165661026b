// Large-sector QR (a9; replaces the per-sector scipy.linalg.qr of abelian.cpp:3123), real float64: two-level blocked
// Householder with the trailing update and the formation of Q on the FP64 tensor pipe.
//
// Reflector j of the outer panel [J0, J0 + nbo) is row j of Wt (n x m, row-major): zeros before column j, its head in
// uhead[j], its tail in Wt[j][j+1:m]; H_j = I - tau_j u_j u_j^T.
//  * qr_factor_outer_panel: the nbo <= 32 panel rows are factored by ONE CTA entirely in shared memory (inner panels of
//    QB_NBI reflectors: unblocked inside, one blocked application to the rest of the outer panel), so no step of the
//    column-by-column chain waits for L2.
//  * qr_block_apply: for a block of rows Y (each row a vector of length m)
//        Y <- Y (I - U^T X U)        X = T (trailing update: H_{J0+nbo-1} ... H_{J0} on every column of A) or T^T (Q formation)
//    as three GEMMs,  D = Y U^T (K = m),  W = D X (as a triangular solve, below),  Y -= W U (K = nbo), with D and the
//    rank-nbo update on DMMA (mma.sync.m8n8k4.f64) from operand chunks that a 3-stage cp.async pipeline keeps in shared memory.
//    T is never formed:  T^-1 = strict_upper(U U^T) + diag(1 / tau)  (the compact-WY recurrence T_{a+1} = [[T_a, -tau T_a z],
//    [0, tau]] inverted), so W = D T is the forward substitution  W[:,a] = tau_a (D[:,a] - sum_{b<a} W[:,b] Z[b][a])  and
//    W = D T^T the backward one; Z = U U^T comes out of the same operand chunks as D.  tau_a = 0 (H_a = I) gives W[:,a] = 0.
#pragma once
#include "common.cuh"

namespace cyb {

constexpr int QB_NBO = 32;        // reflectors per outer panel (at most; fewer when 32 rows do not fit shared memory)
constexpr int QB_NBI = 4;         // reflectors per inner panel
constexpr int QB_RB = 64;         // rows of Y per work block
constexpr int QB_KC = 64;         // columns per staged operand chunk
constexpr int QB_ST = 3;          // pipeline stages
constexpr int QB_THREADS = 512;   // threads per CTA of the blocked kernel
constexpr int QB_LD = QB_KC + 4;  // shared-memory row stride of the chunks: = 4 mod 16 -> conflict-free DMMA fragment loads
constexpr int QB_LDW = QB_NBO + 4;
constexpr int QB_LDZ = QB_NBO + 1;
constexpr size_t QB_STAGE_DOUBLES = (size_t)(QB_NBO + QB_RB) * QB_LD;
// doubles of dynamic shared memory of the block apply: stages of (Uc | Yc), Dsm, Zsm, tau
constexpr size_t QB_SMEM_DOUBLES = QB_ST * QB_STAGE_DOUBLES + (size_t)QB_RB * QB_LDW + (size_t)QB_NBO * QB_LDZ + QB_NBO;
constexpr size_t QB_SMEM_LIMIT = 200 * 1024;   // bytes of dynamic shared memory the kernel may ask for

__device__ __forceinline__ void qb_dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void qb_cp8(double* smem, const double* g)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(g) : "memory");
}
__device__ __forceinline__ void qb_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void qb_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
// optional phase timers (CYB_QR_PROFILE): pt[slot] += time since the previous mark, as seen by the calling thread
struct QbProf {
    unsigned long long* pt;
    unsigned long long last;
    __device__ __forceinline__ void mark(int slot)
    {
        if (!pt) return;
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (slot >= 0) pt[slot] += now - last;
        last = now;
    }
};

// ---- the outer panel, factored in shared memory by the calling CTA (all 1024 threads) -------------------------------
// Ps: (nbo_cap + QB_NBI) x m doubles of shared memory (nbo <= nbo_cap).  On return rows [J0, J0 + nbo) of Wt hold R (columns <= row) and the reflector
// tails, uhead / tau the heads and scalars (global memory, fenced).
__device__ __forceinline__ void qr_factor_outer_panel(double* __restrict__ Wt, double* __restrict__ uhead, double* __restrict__ tau,
                                      int64_t J0, int nbo, int nbo_cap, int64_t m, double* Ps, unsigned long long* pt = nullptr)
{
    QbProf P{pt, 0};
    P.mark(-1);
    __shared__ double s_part[32];
    __shared__ double s_dot[4][8];
    __shared__ double s_head[QB_NBO], s_tau[QB_NBO];
    __shared__ double s_T[QB_NBI][QB_NBI], s_z[QB_NBI][QB_NBI];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nthr = blockDim.x;
    __syncthreads();
    for (int64_t e = tid; e < (int64_t)nbo * m; e += nthr) qb_cp8(Ps + e, Wt + J0 * m + e);   // all in flight at once
    qb_commit();
    qb_wait<0>();
    __syncthreads();
    P.mark(6);
    double* Us = Ps + (int64_t)nbo_cap * m;               // QB_NBI x m: the current inner panel's reflectors, explicit
    for (int i0 = 0; i0 < nbo; i0 += QB_NBI) {
        const int ie = (i0 + QB_NBI < nbo) ? i0 + QB_NBI : nbo;
        for (int aa = i0; aa < ie; ++aa) {
            const int64_t j = J0 + aa;
            double* x = Ps + (int64_t)aa * m;
            double part = 0.0;
            for (int64_t q = j + tid; q < m; q += nthr) part += x[q] * x[q];
            part = warp_sum(part);
            if (lane == 0) s_part[warp] = part;
            __syncthreads();
            double nrm2 = 0.0;
#pragma unroll
            for (int w = 0; w < 32; ++w) nrm2 += w < (nthr >> 5) ? s_part[w] : 0.0;
            const double x0 = x[j];
            const double ax0 = fabs(x0), nrm = sqrt(nrm2);
            const double alpha = (x0 >= 0.0 ? -nrm : nrm);        // -sign(x0) ||x||  (sign(0) = +1)
            const double u0 = x0 - alpha;
            const double uu = nrm2 - ax0 * ax0 + u0 * u0;
            const double tj = (nrm > 0.0 && uu > 0.0) ? 2.0 / uu : 0.0;
            __syncthreads();                                       // x[j] and the partial sums have been read
            if (tid == 0) {
                tau[j] = tj;
                uhead[j] = u0;
                s_head[aa] = u0;
                s_tau[aa] = tj;
                x[j] = tj != 0.0 ? alpha : x0;
            }
            // H_j on the remaining rows of the inner panel: a quarter of the warps per row
            const int wpr = (nthr >> 5) >> 2;                       // warps per row (4 rows at most)
            const int r = warp / wpr, prt = warp % wpr;
            const bool act = aa + 1 + r < ie;
            double* y = Ps + (int64_t)(aa + 1 + r) * m;
            if (act) {
                double dot = (prt == 0 && lane == 0) ? u0 * y[j] : 0.0;
                for (int64_t q = j + 1 + prt * 32 + lane; q < m; q += wpr * 32) dot += x[q] * y[q];
                dot = warp_sum(dot);
                if (lane == 0) s_dot[r][prt] = dot;
            }
            __syncthreads();
            if (act) {
                double dot = 0.0;
#pragma unroll
                for (int w = 0; w < 8; ++w) dot += w < wpr ? s_dot[r][w] : 0.0;
                const double f = tj * dot;
                if (prt == 0 && lane == 0) y[j] -= f * u0;
                for (int64_t q = j + 1 + prt * 32 + lane; q < m; q += wpr * 32) y[q] -= f * x[q];
            }
            __syncthreads();
        }
        P.mark(7);
        if (ie >= nbo) break;
        // the rest of the outer panel takes the inner panel's reflectors at once: y <- (I - U T^T U^T) y.
        // Us: the nb reflectors in explicit form (zeros before the head), so the sweeps below are branch-free
        const int nb = ie - i0;
        const int64_t j0 = J0 + i0;
        for (int64_t e = tid; e < (int64_t)nb * m; e += nthr) {
            const int ab = (int)(e / m);
            const int64_t q = e - (int64_t)ab * m, j = j0 + ab;
            Us[e] = q > j ? Ps[(int64_t)(i0 + ab) * m + q] : (q == j ? s_head[i0 + ab] : 0.0);
        }
        __syncthreads();
        if (warp < nb * nb) {
            const int bb = warp / nb, ab = warp % nb;
            if (bb < ab) {
                double dot = 0.0;
                for (int64_t q = j0 + ab + lane; q < m; q += 32) dot += Us[(int64_t)bb * m + q] * Us[(int64_t)ab * m + q];
                dot = warp_sum(dot);
                if (lane == 0) s_z[bb][ab] = dot;
            }
        }
        __syncthreads();
        if (tid == 0) {
            for (int ab = 0; ab < nb; ++ab) {
                const double ta = s_tau[i0 + ab];
                for (int bb = 0; bb < QB_NBI; ++bb) s_T[bb][ab] = 0.0;
                s_T[ab][ab] = ta;
                for (int bb = 0; bb < ab; ++bb) {                 // T[0:a, a] = -tau_a T[0:a, 0:a] z[0:a, a]
                    double acc = 0.0;
                    for (int cc = bb; cc < ab; ++cc) acc += s_T[bb][cc] * s_z[cc][ab];
                    s_T[bb][ab] = -ta * acc;
                }
            }
        }
        __syncthreads();
        // two rows per warp share every reflector element they read (the sweep is bound by shared-memory bandwidth)
        for (int row = ie + 2 * warp; row < nbo; row += 2 * (nthr >> 5)) {
            double* y0 = Ps + (int64_t)row * m;
            const bool two = row + 1 < nbo;
            double* y1 = two ? y0 + m : y0;
            double d0[QB_NBI], d1[QB_NBI];
#pragma unroll
            for (int ab = 0; ab < QB_NBI; ++ab) d0[ab] = d1[ab] = 0.0;
            for (int64_t q = j0 + lane; q < m; q += 32) {
                const double v0 = y0[q], v1 = y1[q];
#pragma unroll
                for (int ab = 0; ab < QB_NBI; ++ab) {
                    const double u = ab < nb ? Us[(int64_t)ab * m + q] : 0.0;
                    d0[ab] += u * v0;
                    d1[ab] += u * v1;
                }
            }
#pragma unroll
            for (int ab = 0; ab < QB_NBI; ++ab) { d0[ab] = warp_sum(d0[ab]); d1[ab] = warp_sum(d1[ab]); }
            double w0[QB_NBI], w1[QB_NBI];
#pragma unroll
            for (int ab = 0; ab < QB_NBI; ++ab) {
                double a0 = 0.0, a1 = 0.0;
#pragma unroll
                for (int bb = 0; bb < QB_NBI; ++bb)
                    if (bb <= ab) { a0 += s_T[bb][ab] * d0[bb]; a1 += s_T[bb][ab] * d1[bb]; }
                w0[ab] = a0; w1[ab] = a1;
            }
            for (int64_t q = j0 + lane; q < m; q += 32) {
                double v0 = y0[q], v1 = y1[q];
#pragma unroll
                for (int ab = 0; ab < QB_NBI; ++ab) {
                    const double u = ab < nb ? Us[(int64_t)ab * m + q] : 0.0;
                    v0 -= u * w0[ab];
                    v1 -= u * w1[ab];
                }
                y0[q] = v0;
                if (two) y1[q] = v1;
            }
        }
        __syncthreads();
        P.mark(8);
    }
    __syncthreads();
    for (int a = 0; a < nbo; ++a)
        for (int64_t q = J0 + tid; q < m; q += nthr) __stcg(Wt + (J0 + a) * m + q, Ps[(int64_t)a * m + q]);
    __threadfence();
    __syncthreads();
    P.mark(9);
}

// ---- compact-WY application of the outer panel on DMMA ---------------------------------------------------------------
// All 1024 threads of every CTA of the team call this.  The rows [row_begin, row_end) of Y (row stride m) are split into
// one contiguous share per CTA (a multiple of 8 rows, at most QB_RB per pass over the columns).
// Warp roles (QB_THREADS = 512, so that a thread may hold 128 registers: with 200 KB of shared memory there is no L1 left
// and every spilled register is an L2 round trip): warps 0..7 multiply (warp w owns the 8 rows [8w, 8w + 8) of the share:
// one A fragment feeds the four reflector tiles in pass 1, W fragments held in registers feed the eight column tiles in
// pass 2 -- fewer shared-memory reads per DMMA than one 8 x 8 tile per warp), warps 8..11 accumulate Z = U U^T on the
// CTA's first pass, warps 12..15 only issue the cp.async of the 3-stage operand pipeline.
__device__ __forceinline__ void qr_block_apply(double* __restrict__ Y, int64_t row_begin, int64_t row_end, const double* __restrict__ Wt,
                               const double* __restrict__ uhead, const double* __restrict__ tau, int64_t J0, int nbo,
                               int64_t m, bool adj, double* smem, unsigned cta, unsigned ncta,
                               unsigned long long* pt = nullptr)
{
    QbProf P{pt, 0};
    P.mark(-1);
    constexpr int NPROD = 4 * 32, PROD0 = 12 * 32;       // producer threads
    double* Dsm = smem + QB_ST * QB_STAGE_DOUBLES;        // D, then W (QB_RB x QB_LDW)
    double* Zsm = Dsm + (size_t)QB_RB * QB_LDW;           // Z, then T (QB_NBO x QB_LDZ)
    double* taus = Zsm + (size_t)QB_NBO * QB_LDZ;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int64_t qs = (J0 / 4) * 4;                    // first column any reflector of the panel touches (rounded down)
    const int nch = (int)((m - qs + QB_KC - 1) / QB_KC);
    // this CTA's share of the rows
    const int64_t nt = row_end - row_begin;
    int64_t share = ((nt + ncta - 1) / ncta + 7) / 8 * 8;
    const int64_t my_begin = row_begin + (int64_t)cta * share;
    const int64_t my_end = (my_begin + share < row_end) ? my_begin + share : row_end;
    bool have_t = false;
    __syncthreads();
    if (tid < QB_NBO) taus[tid] = tid < nbo ? __ldcg(&tau[J0 + tid]) : 0.0;
    const bool producer = tid >= PROD0;
    const int ptid = tid - PROD0;

    // Producer thread p copies column (p % QB_KC) of rows p / QB_KC, + 6, + 12, ... of a chunk: the source and destination
    // pointers advance by constant strides, so a copy costs a handful of instructions.
    // U chunk: rows = reflectors (zeros before the head, the head from uhead, the tail from Wt), into stage buffer `buf`
    const int pq = ptid & (QB_KC - 1), pr0 = ptid / QB_KC;
    constexpr int PROWS = NPROD / QB_KC;                 // rows covered per sweep of the producer threads (6)
    auto stage_u = [&](double* buf, int64_t q0) {
        const int64_t q = q0 + pq;
        double* dst = buf + pr0 * QB_LD + pq;
        const double* src = Wt + (J0 + pr0) * m + q;
        int64_t j = J0 + pr0;
        for (int a = pr0; a < QB_NBO; a += PROWS, dst += PROWS * QB_LD, src += PROWS * m, j += PROWS) {
            if (a < nbo && q < m && q > j) qb_cp8(dst, src);
            else *dst = (a < nbo && q == j) ? __ldcg(&uhead[j]) : 0.0;
        }
    };
    auto stage_y = [&](double* buf, int64_t q0, int64_t r0, int nr, int nr8) {
        const int64_t q = q0 + pq;
        double* dst = buf + pr0 * QB_LD + pq;
        const double* src = Y + (r0 + pr0) * m + q;
        const bool inq = q < m;
        for (int r = pr0; r < nr8; r += PROWS, dst += PROWS * QB_LD, src += PROWS * m) {
            if (inq && r < nr) qb_cp8(dst, src);
            else *dst = 0.0;
        }
    };

    for (int64_t r0 = my_begin; r0 < my_end; r0 += QB_RB) {
        const int nr = (int)((my_end - r0) < QB_RB ? (my_end - r0) : QB_RB);
        const int nr8 = (nr + 7) / 8 * 8;
        const int rt = warp;
        const bool mul = warp < 8 && rt * 8 < nr8;       // warps 0..7: row tile rt
        const bool do_z = !have_t && warp >= 8 && warp < 12;
        const int zr = warp - 8;
        double* Wsm = smem + QB_NBO * QB_LD;             // W (QB_RB x QB_LD): the Y part of stage 0, unused in pass 2
        double acc[4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = 0.0;
        // ---- pass 1: D = Y U^T (and Z = U U^T once per CTA), 3-stage pipeline over the column chunks ----
        __syncthreads();                                 // the stage buffers and Dsm of the previous pass are free
        if (producer)
            for (int c = 0; c < QB_ST - 1; ++c) {
                if (c < nch) {
                    double* st = smem + (size_t)c * QB_STAGE_DOUBLES;
                    stage_u(st, qs + (int64_t)c * QB_KC);
                    stage_y(st + QB_NBO * QB_LD, qs + (int64_t)c * QB_KC, r0, nr, nr8);
                }
                qb_commit();
            }
        for (int c = 0; c < nch; ++c) {
            if (producer) qb_wait<QB_ST - 2>();
            __syncthreads();                             // chunk c has landed; the buffer of chunk c-1 is free
            if (producer) {
                const int cn = c + QB_ST - 1;
                if (cn < nch) {
                    double* st = smem + (size_t)(cn % QB_ST) * QB_STAGE_DOUBLES;
                    stage_u(st, qs + (int64_t)cn * QB_KC);
                    stage_y(st + QB_NBO * QB_LD, qs + (int64_t)cn * QB_KC, r0, nr, nr8);
                }
                qb_commit();
            } else if (mul || do_z) {
                const double* Uc = smem + (size_t)(c % QB_ST) * QB_STAGE_DOUBLES;
                const double* arow = mul ? Uc + QB_NBO * QB_LD + (rt * 8 + g) * QB_LD + t : Uc + (zr * 8 + g) * QB_LD + t;
                const double* ub = Uc + g * QB_LD + t;
#pragma unroll
                for (int k0 = 0; k0 < QB_KC; k0 += 4) {
                    const double a = arow[k0];
#pragma unroll
                    for (int i = 0; i < 4; ++i) qb_dmma(acc[i][0], acc[i][1], a, ub[i * 8 * QB_LD + k0]);
                }
            }
        }
        if (producer) qb_wait<0>();
        if (mul) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Dsm[(rt * 8 + g) * QB_LDW + i * 8 + 2 * t] = acc[i][0];
                Dsm[(rt * 8 + g) * QB_LDW + i * 8 + 2 * t + 1] = acc[i][1];
            }
        } else if (do_z) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Zsm[(zr * 8 + g) * QB_LDZ + i * 8 + 2 * t] = acc[i][0];
                Zsm[(zr * 8 + g) * QB_LDZ + i * 8 + 2 * t + 1] = acc[i][1];
            }
        }
        __syncthreads();
        P.mark(10);
        // pass 2 needs U only: the producers start its pipeline while warp 0 inverts T
        if (producer)
            for (int c = 0; c < QB_ST - 1; ++c) {
                if (c < nch) stage_u(smem + (size_t)c * QB_STAGE_DOUBLES, qs + (int64_t)c * QB_KC);
                qb_commit();
            }
        // ---- T = (strict_upper(Z) + diag(1/tau))^-1, column c by lane c of warp 0 (in place: column c of Zsm below and on
        //      the diagonal is written only by lane c, the strict upper part of Z that other lanes still read lives in rows
        //      above their diagonal -- so T goes to the LOWER triangle + diagonal as T^T, Z stays in the upper one) ----
        if (!have_t && warp == 0) {
            const int c = lane;
            // x[i] (i <= c) is kept at Zsm[c][i] (row c, left of and on the diagonal)
            if (c < QB_NBO) Zsm[c * QB_LDZ + c] = taus[c];
            __syncwarp();
            for (int i = QB_NBO - 2; i >= 0; --i) {
                // lanes c > i:  x[i] = -tau_i sum_{k=i+1..c} Z[i][k] x[k]
                if (c > i) {
                    double s0 = 0.0, s1 = 0.0;
                    int k = i + 1;
                    for (; k + 1 <= c; k += 2) {
                        s0 += Zsm[i * QB_LDZ + k] * Zsm[c * QB_LDZ + k];
                        s1 += Zsm[i * QB_LDZ + k + 1] * Zsm[c * QB_LDZ + k + 1];
                    }
                    if (k <= c) s0 += Zsm[i * QB_LDZ + k] * Zsm[c * QB_LDZ + k];
                    Zsm[c * QB_LDZ + i] = -taus[i] * (s0 + s1);
                }
                __syncwarp();
            }
        }
        have_t = true;
        __syncthreads();
        // ---- W = D X on DMMA: X[b][a] = T[b][a] (adj) or T[a][b] (Q formation), T[b][a] = Zsm[a][b] for b <= a ----
        if (mul) {
            double w[4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i][0] = w[i][1] = 0.0;
            const double* da = Dsm + (rt * 8 + g) * QB_LDW + t;
#pragma unroll
            for (int k0 = 0; k0 < QB_NBO; k0 += 4) {
                const double a = da[k0];
                const int b = k0 + t;                    // B fragment: X[b = k0 + t][col]
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int col = i * 8 + g;
                    double x;
                    if (adj) x = b <= col ? Zsm[col * QB_LDZ + b] : 0.0;
                    else x = col <= b ? Zsm[b * QB_LDZ + col] : 0.0;
                    qb_dmma(w[i][0], w[i][1], a, x);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Wsm[(rt * 8 + g) * QB_LD + i * 8 + 2 * t] = w[i][0];
                Wsm[(rt * 8 + g) * QB_LD + i * 8 + 2 * t + 1] = w[i][1];
            }
        }
        __syncthreads();
        P.mark(11);
        // ---- pass 2: Y -= W U; a warp keeps its W fragments in registers and sweeps the 8 column tiles of every chunk ----
        double wf[8];
        if (mul) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) wf[kk] = Wsm[(rt * 8 + g) * QB_LD + kk * 4 + t];
        }
        for (int c = 0; c < nch; ++c) {
            const int64_t q0 = qs + (int64_t)c * QB_KC;
            const int yr = rt * 8 + g;
            double yv[8][2];
            if (mul) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int64_t q = q0 + i * 8 + 2 * t;
                    const double* y = Y + (r0 + yr) * m + q;
                    yv[i][0] = (yr < nr && q < m) ? __ldcg(y) : 0.0;
                    yv[i][1] = (yr < nr && q + 1 < m) ? __ldcg(y + 1) : 0.0;
                }
            }
            if (producer) qb_wait<QB_ST - 2>();
            __syncthreads();
            if (producer) {
                const int cn = c + QB_ST - 1;
                if (cn < nch) stage_u(smem + (size_t)(cn % QB_ST) * QB_STAGE_DOUBLES, qs + (int64_t)cn * QB_KC);
                qb_commit();
            } else if (mul) {
                const double* Uc = smem + (size_t)(c % QB_ST) * QB_STAGE_DOUBLES;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    double c0 = 0.0, c1 = 0.0;
                    const double* ub = Uc + t * QB_LD + i * 8 + g;
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) qb_dmma(c0, c1, wf[kk], ub[kk * 4 * QB_LD]);
                    if (yr < nr) {
                        const int64_t q = q0 + i * 8 + 2 * t;
                        double* y = Y + (r0 + yr) * m + q;
                        if (q < m) __stcg(y, yv[i][0] - c0);
                        if (q + 1 < m) __stcg(y + 1, yv[i][1] - c1);
                    }
                }
            }
        }
        if (producer) qb_wait<0>();
        P.mark(12);
    }
    __syncthreads();
}

}  // namespace cyb
