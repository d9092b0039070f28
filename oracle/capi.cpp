// CPU ORACLE (test infrastructure only) — flat C interface for ctypes (tests/, bench.py cpu_baseline, smoke()).
// Also holds the anchor/window schedule of src/main_multi_frame.cpp:109-201 (fit_sequence) and the
// AA -> rotation write-back of Sim3BA.h:481-507 / MultiFrameBA.h:154-172.
#include <omp.h>

#include "oracle.hpp"
#include "resid.hpp"

using namespace orc;

namespace {

// write-back a15: R0 <- R(rootAA) R0 (in place, cumulative), r[0] = R0, r[j] = R(jointAA_j), p = rootT
void writeback_frame(const double* pose /*76*/, double* R0 /*9 in-out*/, double* r /*24*9 or null*/, double* p /*3 or null*/) {
    double Rr[9];
    aa_to_matrix(pose + IDX_RAA, Rr);
    matmul3_cm(Rr, R0, R0);
    if (r) {
        for (int k = 0; k < 9; ++k) r[k] = R0[k];
        for (int j = 1; j < NJ; ++j) aa_to_matrix(pose + idx_jaa(j), r + 9 * j);
    }
    if (p) for (int c = 0; c < 3; ++c) p[c] = pose[IDX_T + c];
}

struct SeqOptions {
    int32_t skip, wsize, overlap, it1, it2, num_shards, nthreads, pad;
    double beta_pose, beta_shape, lambda_temp, beta_shape_lock;
};

}  // namespace

extern "C" {

void* orc_model_create(int V, int J, int S, const double* v_template, const double* shapedirs, const double* J_regressor,
                       const double* weights, const uint32_t* kintree_row0, const double* posedirs, int P) {
    if (J != NJ || S != NS) return nullptr;
    Model* m = new Model();
    m->derive(V, J, S, v_template, shapedirs, J_regressor, weights, kintree_row0, posedirs, P);
    return m;
}
void orc_model_free(void* m) { delete (Model*)m; }
void orc_model_get(void* mp, int32_t* parent, double* J0, double* JSR, double* offset) {
    const Model& M = *(Model*)mp;
    if (parent) for (int j = 0; j < M.J; ++j) parent[j] = M.parent[j];
    if (J0) std::memcpy(J0, M.J0.data(), sizeof(double) * 3 * M.J);
    if (JSR) std::memcpy(JSR, M.JSR.data(), sizeof(double) * 3 * M.J * M.S);
    if (offset) rest_offsets(M, offset);
}

void orc_forward(void* mp, int B, const double* w, const double* p, const double* r, double* cloud, double* jointPos,
                 double* jointTrans, int nthreads) {
    const Model& M = *(Model*)mp;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(static)
    for (int b = 0; b < B; ++b)
        M.forward(w + (size_t)b * M.S, p + 3 * (size_t)b, r + 9 * (size_t)M.J * b, cloud ? cloud + 3 * (size_t)M.V * b : nullptr,
                  jointPos ? jointPos + 3 * (size_t)M.J * b : nullptr, jointTrans ? jointTrans + 12 * (size_t)M.J * b : nullptr);
}

void* orc_gmm_load(const char* path, int* status) {
    GMM* g = new GMM();
    const int st = g->load_text(path);
    if (status) *status = st;
    return g;
}
void* orc_gmm_create(int nc, int nd, const double* weights, const double* means, const double* covs, int* status) {
    GMM* g = new GMM();
    const int st = g->build(nc, nd, weights, means, covs);
    if (status) *status = st;
    return g;
}
void orc_gmm_free(void* g) { delete (GMM*)g; }
void orc_gmm_dims(void* gp, int* nc, int* nd) { *nc = ((GMM*)gp)->nComps; *nd = ((GMM*)gp)->nDims; }
void orc_gmm_get(void* gp, double* weights, double* means, double* covs, double* prec_cho, double* consts_log) {
    const GMM& g = *(GMM*)gp;
    const size_t nn = (size_t)g.nDims * g.nDims;
    for (int i = 0; i < g.nComps; ++i) {
        if (weights) weights[i] = g.weight[i];
        if (consts_log) consts_log[i] = g.consts_log[i];
        if (covs) std::memcpy(covs + i * nn, g.cov[i].data(), sizeof(double) * nn);
        if (prec_cho) std::memcpy(prec_cho + i * nn, g.prec_cho[i].data(), sizeof(double) * nn);
    }
    if (means) std::memcpy(means, g.mean.data(), sizeof(double) * (size_t)g.nComps * g.nDims);
}
int orc_gmm_residual(void* gp, const double* x, double* out) { return ((GMM*)gp)->residual(x, out); }

void orc_aa_to_matrix(const double* aa, double* R) { aa_to_matrix(aa, R); }

// residual + Jacobian of one reprojection block, three ways (for the oracle self-tests)
void orc_reproj_check(void* mp, int with_shape, int jid, double u, double v, const double* cam, const double* R0,
                      const double* x86, double* r_plain, double* r_jet, double* J_jet, double* r_ana, double* J_ana) {
    const Model& M = *(Model*)mp;
    Skeleton sk;
    sk.J = M.J; sk.S = M.S;
    for (int j = 0; j < M.J; ++j) sk.parent[j] = M.parent[j];
    rest_offsets(M, sk.offset);
    sk.JSR = with_shape ? M.JSR.data() : nullptr;
    reproj_residual<double>(sk, jid, u, v, cam, R0, x86, r_plain);
    std::vector<Jet<NFULL>> xj(NFULL);
    for (int k = 0; k < NFULL; ++k) xj[k] = Jet<NFULL>(x86[k], k);
    Jet<NFULL> res[2];
    reproj_residual<Jet<NFULL>>(sk, jid, u, v, cam, R0, xj.data(), res);
    for (int i = 0; i < 2; ++i) { r_jet[i] = res[i].a; for (int k = 0; k < NFULL; ++k) J_jet[i * NFULL + k] = res[i].v[k]; }
    reproj_analytic(sk, jid, u, v, cam, R0, x86, r_ana, J_ana);
}

// B independent single-frame fits (the frame loop of src/main_single_frame.cpp:192), OpenMP over frames.
// trace: B x trace_stride IterRecord (or null).  Returns number of usable fits.
int orc_fit_single_batch(void* mp, void* gp, const FitOptions* o, int B, const KP* kps, const int32_t* nkp, const double* R0,
                         double* sim3, double* w, double* jointAA, uint8_t* usable, IterRecord* trace, int trace_stride,
                         int32_t* niters, int32_t* term, int nthreads) {
    const Model& M = *(Model*)mp;
    const GMM* g = (const GMM*)gp;
    int ok = 0;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(dynamic, 1) reduction(+ : ok)
    for (int b = 0; b < B; ++b) {
        std::vector<IterRecord> tr;
        LMSummary s;
        const bool u = fit_single(M, g, *o, kps + (size_t)b * MAXKP, nkp[b], R0 + 9 * (size_t)b, sim3 + 7 * (size_t)b,
                                  w + (size_t)NS * b, jointAA + 3 * (size_t)NJ * b, trace ? &tr : nullptr, &s);
        if (usable) usable[b] = u ? 1 : 0;
        if (niters) niters[b] = s.num_iterations;
        if (term) term[b] = s.termination;
        if (trace)
            for (int i = 0; i < (int)tr.size() && i < trace_stride; ++i) trace[(size_t)b * trace_stride + i] = tr[i];
        ok += u ? 1 : 0;
    }
    return ok;
}

int orc_fit_window(void* mp, const FitOptions* o, int F, const KP* kps, const int32_t* nkp, const double* R0, double* poses,
                   double* w, IterRecord* trace, int trace_cap, int32_t* niters, int32_t* term) {
    const Model& M = *(Model*)mp;
    std::vector<IterRecord> tr;
    LMSummary s;
    const bool u = fit_window(M, *o, F, kps, nkp, R0, poses, w, trace ? &tr : nullptr, &s);
    if (niters) *niters = s.num_iterations;
    if (term) *term = s.termination;
    if (trace)
        for (int i = 0; i < (int)tr.size() && i < trace_cap; ++i) trace[i] = tr[i];
    return u ? 1 : 0;
}

void orc_writeback(int F, const double* poses, double* R0, double* r, double* p) {
    for (int f = 0; f < F; ++f)
        writeback_frame(poses + (size_t)f * NPF, R0 + 9 * (size_t)f, r ? r + 9 * (size_t)NJ * f : nullptr, p ? p + 3 * (size_t)f : nullptr);
}

// 3dba_multi schedule (src/main_multi_frame.cpp:88-201).  State arrays (all N frames):
//   poses N x 76 (in: initial FramePoseParams; out: refined), R0 N x 9 (in: avatars[i]->r[0]; out: after cumulative write-back),
//   W N x 10 (in: avatars[i]->w; out), r N x 24 x 9 and p N x 3 (out).
// num_shards == 1 reproduces the reference's serial chain exactly; num_shards > 1 is the sharded schedule of DESIGN.md
// (pass 1 per shard from unrefined halo, halo hand-off, one fix-up re-solve of each shard's first window).
// win_iters / win_term (optional): per executed solve, order = [stage1, pass-1 windows in index order, fix-ups in shard order].
int orc_fit_sequence(void* mp, const FitOptions* base, const SeqOptions* so, int N, const KP* kps, const int32_t* nkp, double* poses,
                     double* R0, double* W, double* r, double* p, int32_t* win_iters, int32_t* win_term) {
    const Model& M = *(Model*)mp;
    int nfail = 0, rec = 0;
    auto record = [&](int slot, const LMSummary& s) { if (win_iters) win_iters[slot] = s.num_iterations; if (win_term) win_term[slot] = s.termination; };
    // ---- stage 1: anchors ----
    std::vector<int> anchors;
    for (int i = 0; i < N; i += so->skip) anchors.push_back(i);
    {
        const int A = (int)anchors.size();
        std::vector<KP> k((size_t)A * MAXKP);
        std::vector<int> nk(A);
        std::vector<double> ps((size_t)A * NPF), r0((size_t)A * 9);
        for (int a = 0; a < A; ++a) {
            const int id = anchors[a];
            std::memcpy(&k[(size_t)a * MAXKP], kps + (size_t)id * MAXKP, sizeof(KP) * MAXKP);
            nk[a] = nkp[id];
            std::memcpy(&ps[(size_t)a * NPF], poses + (size_t)id * NPF, sizeof(double) * NPF);   // anchorPos is a COPY (:117)
            std::memcpy(&r0[(size_t)a * 9], R0 + 9 * (size_t)id, sizeof(double) * 9);
        }
        FitOptions o = *base;
        o.beta_pose = so->beta_pose; o.beta_shape = so->beta_shape; o.lambda_temp = so->lambda_temp; o.max_iters = so->it1;
        o.linear_solver = LS_BLOCK_TRIDIAG;
        LMSummary s;
        if (!fit_window(M, o, A, k.data(), nk.data(), r0.data(), ps.data(), W /*avatars.front()->w*/, nullptr, &s)) ++nfail;
        record(rec++, s);
        for (int a = 0; a < A; ++a) {   // write-back into the avatars only; poses[] keep their initial values
            const int id = anchors[a];
            writeback_frame(&ps[(size_t)a * NPF], R0 + 9 * (size_t)id, r + 9 * (size_t)NJ * id, p + 3 * (size_t)id);
        }
    }
    for (int i = 1; i < N; ++i) std::memcpy(W + (size_t)NS * i, W, sizeof(double) * NS);   // share shape (:154)

    // ---- stage 2: sliding windows ----
    const int stride = so->wsize - so->overlap;
    std::vector<int> starts;
    for (int s = 0; s < N; s += stride) starts.push_back(s);
    const int K = (int)starts.size();
    const int G = std::max(1, std::min<int>(so->num_shards, K));
    FitOptions o = *base;
    o.beta_pose = so->beta_pose; o.beta_shape = so->beta_shape_lock; o.lambda_temp = so->lambda_temp; o.max_iters = so->it2;
    o.linear_solver = LS_BLOCK_TRIDIAG;
    auto solve_window = [&](int s, double* Ps, double* R0s, double* Ws, double* rs, double* pp, LMSummary* sm) {
        // arrays are indexed by absolute frame
        const int e = std::min(s + so->wsize, N);
        const bool ok = fit_window(M, o, e - s, kps + (size_t)s * MAXKP, nkp + s, R0s + 9 * (size_t)s, Ps + (size_t)s * NPF,
                                   Ws + (size_t)NS * s /*window-front avatar's w*/, nullptr, sm);
        for (int f = s; f < e; ++f) writeback_frame(Ps + (size_t)f * NPF, R0s + 9 * (size_t)f, rs + 9 * (size_t)NJ * f, pp + 3 * (size_t)f);
        return ok;
    };
    if (G == 1) {
        for (int kidx = 0; kidx < K; ++kidx) {
            LMSummary s;
            if (!solve_window(starts[kidx], poses, R0, W, r, p, &s)) ++nfail;
            record(rec + kidx, s);
        }
        return nfail;
    }
    // sharded schedule
    std::vector<int> kb(G + 1);
    for (int g = 0; g <= G; ++g) kb[g] = (int)((long long)g * K / G);
    const std::vector<double> poses0(poses, poses + (size_t)N * NPF), R00(R0, R0 + (size_t)N * 9), W0(W, W + (size_t)N * NS);
    // pass 1: every shard on a private copy of the post-stage-1 state
    std::vector<std::vector<double>> sp(G), sr0(G), sw(G), srr(G), spp(G);
    std::vector<LMSummary> sums(K);
    std::vector<int> fails(G, 0);
#pragma omp parallel for num_threads(so->nthreads > 0 ? so->nthreads : 1) schedule(dynamic, 1)
    for (int g = 0; g < G; ++g) {
        sp[g] = poses0; sr0[g] = R00; sw[g] = W0;
        srr[g].assign((size_t)N * NJ * 9, 0.0); spp[g].assign((size_t)N * 3, 0.0);
        for (int kidx = kb[g]; kidx < kb[g + 1]; ++kidx)
            if (!solve_window(starts[kidx], sp[g].data(), sr0[g].data(), sw[g].data(), srr[g].data(), spp[g].data(), &sums[kidx])) ++fails[g];
    }
    for (int kidx = 0; kidx < K; ++kidx) record(rec + kidx, sums[kidx]);
    rec += K;
    auto copy_frames = [&](int g, int f0, int f1) {
        for (int f = f0; f < f1; ++f) {
            std::memcpy(poses + (size_t)f * NPF, &sp[g][(size_t)f * NPF], sizeof(double) * NPF);
            std::memcpy(R0 + 9 * (size_t)f, &sr0[g][9 * (size_t)f], sizeof(double) * 9);
            std::memcpy(W + (size_t)NS * f, &sw[g][(size_t)NS * f], sizeof(double) * NS);
            std::memcpy(r + 9 * (size_t)NJ * f, &srr[g][9 * (size_t)NJ * f], sizeof(double) * 9 * NJ);
            std::memcpy(p + 3 * (size_t)f, &spp[g][3 * (size_t)f], sizeof(double) * 3);
        }
    };
    // ownership: shard g owns frames [start(kb[g]), start(kb[g+1])) (last shard: to N)
    for (int g = 0; g < G; ++g) copy_frames(g, starts[kb[g]], g + 1 < G ? starts[kb[g + 1]] : N);
    // pass 2: halo hand-off + fix-up of each shard's first window (serial here; independent across shards)
    std::vector<std::vector<double>> fp(G), fr0(G), fw(G), frr(G), fpp(G);
    std::vector<LMSummary> fsum(G);
#pragma omp parallel for num_threads(so->nthreads > 0 ? so->nthreads : 1) schedule(dynamic, 1)
    for (int g = 1; g < G; ++g) {
        const int s = starts[kb[g]], e = std::min(s + so->wsize, N), h = std::min(s + so->overlap, e);
        fp[g] = poses0; fr0[g] = R00; fw[g] = W0;
        frr[g].assign((size_t)N * NJ * 9, 0.0); fpp[g].assign((size_t)N * 3, 0.0);
        for (int f = s; f < h; ++f) {   // halo = refined overlap frames as left by shard g-1's last window
            std::memcpy(&fp[g][(size_t)f * NPF], &sp[g - 1][(size_t)f * NPF], sizeof(double) * NPF);
            std::memcpy(&fr0[g][9 * (size_t)f], &sr0[g - 1][9 * (size_t)f], sizeof(double) * 9);
        }
        if (!solve_window(s, fp[g].data(), fr0[g].data(), fw[g].data(), frr[g].data(), fpp[g].data(), &fsum[g])) ++fails[g];
    }
    for (int g = 1; g < G; ++g) {
        record(rec++, fsum[g]);
        const int s = starts[kb[g]], e = std::min(s + so->wsize, N);
        const int keep_to = (kb[g + 1] - kb[g] > 1) ? std::min(s + stride, e) : e;   // frames the shard's 2nd window revisits keep pass-1 values
        std::swap(sp[g], fp[g]); std::swap(sr0[g], fr0[g]); std::swap(sw[g], fw[g]); std::swap(srr[g], frr[g]); std::swap(spp[g], fpp[g]);
        copy_frames(g, s, keep_to);
    }
    for (int g = 0; g < G; ++g) nfail += fails[g];
    return nfail;
}

// vertex-level ICP term (icp.cpp): B bodies, each with ncorr[b] correspondences starting at cptr[b]; data cloud of body b
// starts at point dptr[b].  Outputs per body: cost, grad (85); resid / jac only for B == 1 style calls (may be null).
void orc_icp_eval(void* mp, int B, const double* w, const double* p, const double* q, const int64_t* cptr, const int32_t* mpt,
                  const int32_t* dpt, const int64_t* dptr, const double* data, int shape_enabled, double* resid, double* jac,
                  double* cost, double* grad, int nthreads) {
    const Model& M = *(Model*)mp;
    const int NC = 3 + 3 * M.J + M.S;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        const int64_t c0 = cptr[b];
        const int n = (int)(cptr[b + 1] - c0);
        icp_eval(M, w + (size_t)b * M.S, p + 3 * (size_t)b, q + 4 * (size_t)M.J * b, n, mpt + c0, dpt + c0, data + 3 * dptr[b],
                 shape_enabled, resid ? resid + 3 * c0 : nullptr, jac ? jac + 3 * c0 * NC : nullptr, cost + b, grad + (size_t)b * NC);
    }
}
void orc_icp_priors(void* gp, int B, double beta_pose, double beta_shape, const double* totals, const double* w, const double* q,
                    double* cost, double* grad) {
    for (int b = 0; b < B; ++b) icp_priors((const GMM*)gp, beta_pose, beta_shape, totals[b], NS, w + (size_t)NS * b, q + 4 * (size_t)NJ * b, cost + b, grad + (size_t)(3 + 3 * NJ + NS) * b);
}
void orc_icp_visible(int B, int V, int F, const uint32_t* faces, const double* cloud, uint8_t* visible) {
    for (int b = 0; b < B; ++b) icp_visible(V, F, faces, cloud + 3 * (size_t)V * b, visible + (size_t)V * b);
}
void orc_icp_nearest(int B, int V, const double* cloud, const int32_t* model_part, const uint8_t* visible, const int64_t* dptr,
                     const double* data, const int32_t* data_part, int32_t* out, int nthreads) {
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b)
        icp_nearest(V, cloud + 3 * (size_t)V * b, model_part, visible ? visible + (size_t)V * b : nullptr, (int)(dptr[b + 1] - dptr[b]),
                    data + 3 * dptr[b], data_part + dptr[b], out + dptr[b]);
}
// the BFGS / Wolfe minimiser on the ICP objective (icp_solve.cpp), body by body (OpenMP over bodies).  trace: B x trace_cap records
// of 7 doubles (cost, cost_change, gradient_max_norm, step_norm, step_size, line-search evaluations, line-search iterations);
// summary: B x 6 doubles (initial cost, final cost, iterations, evaluations, restarts, termination)
void orc_icp_minimize(void* mp, void* gp, int B, double beta_pose, double beta_shape, const int64_t* cptr, const int32_t* mpt,
                      const int32_t* dpt, const int64_t* dptr, const double* data, int shape_enabled, int max_iterations,
                      double function_tolerance, double gradient_tolerance, double parameter_tolerance, double* w, double* p, double* q,
                      double* trace, int trace_cap, int32_t* nrec, double* summary, int nthreads) {
    const Model& M = *(Model*)mp;
    IcpSolveOptions o;
    o.max_iterations = max_iterations; o.function_tolerance = function_tolerance; o.gradient_tolerance = gradient_tolerance;
    o.parameter_tolerance = parameter_tolerance;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        const int64_t c0 = cptr[b];
        std::vector<IcpSolveRecord> tr;
        IcpSolveSummary sm;
        icp_minimize(M, (const GMM*)gp, beta_pose, beta_shape, (int)(cptr[b + 1] - c0), mpt + c0, dpt + c0, data + 3 * dptr[b], shape_enabled,
                     o, w + (size_t)b * M.S, p + 3 * (size_t)b, q + 4 * (size_t)M.J * b, &tr, &sm);
        const int n = std::min<int>((int)tr.size(), trace_cap);
        if (nrec) nrec[b] = (int)tr.size();
        for (int i = 0; i < n && trace; ++i) {
            double* t = trace + ((size_t)b * trace_cap + i) * 7;
            t[0] = tr[i].cost; t[1] = tr[i].cost_change; t[2] = tr[i].gradient_max_norm; t[3] = tr[i].step_norm; t[4] = tr[i].step_size;
            t[5] = tr[i].line_search_evals; t[6] = tr[i].line_search_iters;
        }
        double* sv = summary + (size_t)b * 6;
        sv[0] = sm.initial_cost; sv[1] = sm.final_cost; sv[2] = sm.iterations; sv[3] = sm.evaluations; sv[4] = sm.restarts; sv[5] = sm.termination;
    }
}
void orc_quat_plus(int n, const double* x, const double* d, double* out) {
    for (int i = 0; i < n; ++i) quat_plus(x + 4 * i, d + 3 * i, out + 4 * i);
}

int orc_sizeof_iter_record() { return (int)sizeof(IterRecord); }
int orc_sizeof_fit_options() { return (int)sizeof(FitOptions); }
int orc_sizeof_kp() { return (int)sizeof(KP); }

}  // extern "C"
