// The anchor / sliding-window schedule of 3dba_multi (src/main_multi_frame.cpp:109-201) on top of the multi-frame LM kernel.
//
//   stage 1  anchors {0, skip, 2 skip, ...}: ONE OptimizeMultiFrame over all of them (max_iters = it1, beta_shape).  anchorPos is
//            a copy (:117), so the optimised per-frame parameters are discarded; what persists is the write-back into the
//            avatars (r, p; MultiFrameBA.h:154-174) and avatars[0]->w, which every avatar then receives (:154).
//   stage 2  windows [s, min(s + wsize, N)), s += wsize - overlap, solved one after the other (:165-201) with beta_shape =
//            beta_shape_lock (1e5) and max_iters = it2 (60); each starts from the refined overlap frames AND their already
//            rotated R0 left by its predecessor (reference behaviour); the shared block is the window-front avatar's w.
//
// num_shards == 1 is that serial chain, exactly.  num_shards = G > 1 is the sharded schedule: the window
// list is cut into G contiguous runs; every run is solved from the post-stage-1 state on a private copy (pass 1), then the first
// window of every run but the first is re-solved starting from the overlap frames its left neighbour refined (pass 2: the halo
// hand-off — OVERLAP x (76 + 9) doubles per edge, what NCCL send/recv carries between GPUs).  Independent sequences and the
// shards of one sequence are batched into the same launches: one CTA per (sequence, shard) chain per wave.
#include <algorithm>
#include <atomic>

#include "lm_device.cuh"

namespace b3d {

// dst[i] = src[idx[i]] for rows of `w` doubles (gather) or dst[idx[i]] = src[i] (scatter)
__global__ void rows_gather(int n, const int32_t* __restrict__ idx, const double* __restrict__ src, double* __restrict__ dst, int w, int scatter) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (size_t)n * w) return;
    const int i = (int)(t / w), c = (int)(t % w);
    if (scatter) dst[(size_t)idx[i] * w + c] = src[t];
    else dst[t] = src[(size_t)idx[i] * w + c];
}
// W[f] = W[seq_first(f)]
__global__ void share_shape(int N, const int32_t* __restrict__ front_of, double* __restrict__ W) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= N * NS) return;
    const int f = t / NS, s = t % NS;
    const int q = front_of[f];
    if (q != f) W[t] = W[(size_t)q * NS + s];
}

static std::atomic<int> g_seq_waves{0};     // debug hook b3d_debug_seq_waves: 1 = stage 2 as one launch per wave (the round-1 schedule)

struct DevBuf {
    char* p = nullptr;
    ~DevBuf() { cudaFree(p); }
    int alloc(size_t bytes) { B3D_CUDA(cudaMalloc(&p, bytes ? bytes : 1)); return B3D_OK; }
};

static int gather(cudaStream_t st, int n, const int32_t* idx, const void* src, void* dst, int w_doubles, bool scatter) {
    if (n <= 0) return B3D_OK;
    const size_t tot = (size_t)n * w_doubles;
    rows_gather<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(n, idx, (const double*)src, (double*)dst, w_doubles, scatter ? 1 : 0);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_debug_seq_waves(int on) { g_seq_waves.store(on); return 0; }

int b3d_fit_sequences(const b3d_model* m, const b3d_fit_opts* base, const b3d_seq_opts* so, int32_t nseq, const int32_t* seq_first,
                      const int32_t* seq_len, int32_t N, const b3d_kp* kps, const int32_t* nkp, double* poses, double* R0, double* W,
                      double* r, double* p, int32_t* solve_iters, int32_t* solve_term, int32_t* nfail_out) {
    if (!m || !base || !so || !seq_first || !seq_len || !kps || !nkp || !poses || !R0 || !W) return set_error(B3D_EINVAL, "b3d_fit_sequences: null argument");
    if (nseq <= 0 || N <= 0) return set_error(B3D_EINVAL, "b3d_fit_sequences: empty input");
    if (so->skip <= 0 || so->wsize <= 0 || so->overlap < 0 || so->overlap >= so->wsize) return set_error(B3D_EINVAL, "b3d_fit_sequences: bad schedule");
    for (int q = 0; q < nseq; ++q)
        if (seq_first[q] < 0 || seq_len[q] <= 0 || seq_first[q] + seq_len[q] > N) return set_error(B3D_EINVAL, "b3d_fit_sequences: sequence %d out of range", q);
    if (int rc = require_device()) return rc;
    if (int rc = validate_keypoints(kps, nkp, (size_t)N, "b3d_fit_sequences")) return rc;
    const int stride = so->wsize - so->overlap;
    const int G = std::max(1, so->num_shards);
    const size_t n = (size_t)N;
    cudaStream_t st = nullptr;

    // ---------------- host-side schedule ----------------
    struct Chain { int q, g, k0, k1; };                       // windows [k0, k1) of sequence q form shard g
    std::vector<std::vector<int>> starts(nseq);
    std::vector<int32_t> anchor_idx, anchor_first(nseq), anchor_cnt(nseq), front_of(N);
    for (int f = 0; f < N; ++f) front_of[f] = f;
    std::vector<Chain> chains;
    int max_anchor = 0;
    for (int q = 0; q < nseq; ++q) {
        const int f0 = seq_first[q], L = seq_len[q];
        anchor_first[q] = (int32_t)anchor_idx.size();
        for (int i = 0; i < L; i += so->skip) anchor_idx.push_back(f0 + i);
        anchor_cnt[q] = (int32_t)anchor_idx.size() - anchor_first[q];
        max_anchor = std::max(max_anchor, (int)anchor_cnt[q]);
        for (int f = 0; f < L; ++f) front_of[f0 + f] = f0;
        for (int s = 0; s < L; s += stride) starts[q].push_back(s);
        const int K = (int)starts[q].size(), Gq = std::min(G, K);
        for (int g = 0; g < Gq; ++g) chains.push_back({q, g, (int)((long long)g * K / Gq), (int)((long long)(g + 1) * K / Gq)});
    }
    const int TA = (int)anchor_idx.size(), NC = (int)chains.size();
    const bool sharded = G > 1;
    int waves = 0;
    for (const Chain& c : chains) waves = std::max(waves, c.k1 - c.k0);

    // problem table: [stage 1: nseq][pass 1: wave-major][pass 2: chains with g > 0]
    std::vector<WinProblem> probs;
    std::vector<int> wave_off(waves + 1, 0);
    // private state copies: chain c works on copy c (rows [c * N, (c + 1) * N)) when sharded, on the main arrays otherwise
    auto base_row = [&](int c) { return sharded ? (size_t)c * n : (size_t)0; };
    size_t ws_need = 0;
    {
        size_t off = 0;
        for (int q = 0; q < nseq; ++q) { probs.push_back({anchor_first[q], anchor_cnt[q], (int64_t)off}); off += win_ws_doubles(anchor_cnt[q]); }
        ws_need = off;
    }
    // every chain owns one workspace slot for all its windows: the chains run unsynchronised (chain mode of fit_window_kernel)
    const size_t slot = win_ws_doubles(so->wsize);
    ws_need = std::max(ws_need, slot * (size_t)NC);
    std::vector<std::vector<int>> wave_chain(waves);
    std::vector<int32_t> order((size_t)NC * std::max(waves, 1), -1);      // order[c][i] = problem of chain c's i-th window
    for (int i = 0; i < waves; ++i) {
        wave_off[i] = (int)probs.size();
        for (int c = 0; c < NC; ++c) {
            const Chain& ch = chains[c];
            if (ch.k0 + i >= ch.k1) continue;
            const int s = starts[ch.q][ch.k0 + i], e = std::min(s + so->wsize, (int)seq_len[ch.q]);
            order[(size_t)c * waves + i] = (int32_t)probs.size();
            probs.push_back({(int32_t)(base_row(c) + seq_first[ch.q] + s), e - s, (int64_t)(slot * (size_t)c)});
            wave_chain[i].push_back(c);
        }
    }
    wave_off[waves] = (int)probs.size();
    const int fix_off = (int)probs.size();
    std::vector<int> fix_chain;
    if (sharded) {
        for (int c = 0; c < NC; ++c) {
            const Chain& ch = chains[c];
            if (ch.g == 0) continue;
            const int s = starts[ch.q][ch.k0], e = std::min(s + so->wsize, (int)seq_len[ch.q]);
            probs.push_back({(int32_t)(base_row(c) + seq_first[ch.q] + s), e - s, (int64_t)(slot * (size_t)c)});     // the fix-up of chain c takes chain c's slot
            fix_chain.push_back(c);
        }
    }
    const int NP = (int)probs.size();

    // ---------------- device state ----------------
    const size_t copies = sharded ? (size_t)NC : 1;
    DevBuf d_kps, d_nkp, d_poses, d_R0, d_W, d_r, d_p, d_orig, d_fix, d_anchor, d_idx, d_front, d_probs, d_fc, d_ws, d_stat;
    const size_t row_state = NPF + 9 + NS;   // poses, R0, W per frame
    if (int rc = d_kps.alloc(sizeof(b3d_kp) * n * MAXKP * copies)) return rc;
    if (int rc = d_nkp.alloc(4 * n * copies)) return rc;
    if (int rc = d_poses.alloc(8 * n * NPF * copies)) return rc;
    if (int rc = d_R0.alloc(8 * n * 9 * copies)) return rc;
    if (int rc = d_W.alloc(8 * n * NS * copies)) return rc;
    if (int rc = d_r.alloc(8 * n * NJ * 9 * copies)) return rc;
    if (int rc = d_p.alloc(8 * n * 3 * copies)) return rc;
    if (int rc = d_orig.alloc(sharded ? 8 * n * row_state : 0)) return rc;
    if (int rc = d_fix.alloc(sharded ? 8 * n * (row_state + NJ * 9 + 3) * (size_t)fix_chain.size() : 0)) return rc;
    const size_t a_bytes = (size_t)TA * (sizeof(b3d_kp) * MAXKP + 8 + 8 * (NPF + 9 + NS + NJ * 9 + 3));
    if (int rc = d_anchor.alloc(a_bytes)) return rc;
    if (int rc = d_idx.alloc(4 * (size_t)TA)) return rc;
    if (int rc = d_front.alloc(4 * n)) return rc;
    if (int rc = d_probs.alloc(sizeof(WinProblem) * (size_t)NP)) return rc;
    if (int rc = d_fc.alloc(2 * sizeof(FitConst))) return rc;
    if (int rc = d_ws.alloc(8 * ws_need)) return rc;
    if (int rc = d_stat.alloc((size_t)NP * 12)) return rc;
    b3d_kp* kps_d = (b3d_kp*)d_kps.p; int32_t* nkp_d = (int32_t*)d_nkp.p;
    double *poses_d = (double*)d_poses.p, *R0_d = (double*)d_R0.p, *W_d = (double*)d_W.p, *r_d = (double*)d_r.p, *p_d = (double*)d_p.p;
    uint8_t* us_d = (uint8_t*)d_stat.p; int32_t* ni_d = (int32_t*)(d_stat.p + (((size_t)NP + 3) / 4) * 4); int32_t* te_d = ni_d + NP;

    b3d_fit_opts o1 = *base, o2 = *base;
    o1.beta_pose = so->beta_pose; o1.beta_shape = so->beta_shape; o1.lambda_temp = so->lambda_temp; o1.max_iters = so->it1;
    o2.beta_pose = so->beta_pose; o2.beta_shape = so->beta_shape_lock; o2.lambda_temp = so->lambda_temp; o2.max_iters = so->it2;
    FitConst fc[2];
    if (int rc = build_fit_const(m, nullptr, &o1, true, &fc[0])) return rc;
    if (int rc = build_fit_const(m, nullptr, &o2, true, &fc[1])) return rc;
    const FitConst* fc_d = (const FitConst*)d_fc.p;
    const WinProblem* probs_d = (const WinProblem*)d_probs.p;
    B3D_CUDA(cudaMemcpy(d_fc.p, fc, sizeof(fc), cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(d_probs.p, probs.data(), sizeof(WinProblem) * (size_t)NP, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(d_idx.p, anchor_idx.data(), 4 * (size_t)TA, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(d_front.p, front_of.data(), 4 * n, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(kps_d, kps, sizeof(b3d_kp) * n * MAXKP, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(nkp_d, nkp, 4 * n, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(poses_d, poses, 8 * n * NPF, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(R0_d, R0, 8 * n * 9, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(W_d, W, 8 * n * NS, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemset(r_d, 0, 8 * n * NJ * 9 * copies));
    B3D_CUDA(cudaMemset(p_d, 0, 8 * n * 3 * copies));
    const int32_t* idx_d = (const int32_t*)d_idx.p;

    // ---------------- stage 1: anchors ----------------
    {
        char* a = d_anchor.p;
        b3d_kp* a_kps = (b3d_kp*)a; a += sizeof(b3d_kp) * MAXKP * (size_t)TA;
        double* a_poses = (double*)a; a += 8 * (size_t)TA * NPF;
        double* a_R0 = (double*)a; a += 8 * (size_t)TA * 9;
        double* a_W = (double*)a; a += 8 * (size_t)TA * NS;
        double* a_r = (double*)a; a += 8 * (size_t)TA * NJ * 9;
        double* a_p = (double*)a; a += 8 * (size_t)TA * 3;
        int32_t* a_nkp = (int32_t*)a;
        static_assert(sizeof(b3d_kp) % 8 == 0, "keypoints are copied as rows of doubles");
        if (int rc = gather(st, TA, idx_d, kps_d, a_kps, (int)(sizeof(b3d_kp) * MAXKP / 8), false)) return rc;
        if (int rc = gather(st, TA, idx_d, poses_d, a_poses, NPF, false)) return rc;            // anchorPos is a COPY (:117)
        if (int rc = gather(st, TA, idx_d, R0_d, a_R0, 9, false)) return rc;
        if (int rc = gather(st, TA, idx_d, W_d, a_W, NS, false)) return rc;
        std::vector<int32_t> a_nk(TA);
        for (int i = 0; i < TA; ++i) a_nk[i] = nkp[anchor_idx[i]];
        B3D_CUDA(cudaMemcpyAsync(a_nkp, a_nk.data(), 4 * (size_t)TA, cudaMemcpyHostToDevice, st));
        B3D_CUDA(cudaStreamSynchronize(st));
        if (int rc = win_launch(fc_d, probs_d, nseq, a_kps, a_nkp, a_R0, a_poses, a_W, us_d, nullptr, 0, ni_d, te_d, (double*)d_ws.p, st, probs.data())) return rc;   // long anchor chains: grid kernel
        if (int rc = b3d_writeback_dev(TA, a_poses, a_R0, a_r, a_p, st)) return rc;               // into the avatars only
        if (int rc = gather(st, TA, idx_d, a_R0, R0_d, 9, true)) return rc;
        if (int rc = gather(st, TA, idx_d, a_r, r_d, NJ * 9, true)) return rc;
        if (int rc = gather(st, TA, idx_d, a_p, p_d, 3, true)) return rc;
        if (int rc = gather(st, TA, idx_d, a_W, W_d, NS, true)) return rc;    // only the front anchor's row changed (the shared block)
        share_shape<<<(unsigned)((n * NS + 255) / 256), 256, 0, st>>>(N, (const int32_t*)d_front.p, W_d);   // (:154)
        B3D_LAUNCH_CHECK();
    }

    // ---------------- stage 2 ----------------
    auto writeback_problem = [&](const WinProblem& pr, double* po, double* r0, double* rr, double* pp) -> int {
        const size_t f = (size_t)pr.first;
        return b3d_writeback_dev(pr.count, po + f * NPF, r0 + f * 9, rr + f * NJ * 9, pp + f * 3, st);
    };
    if (sharded) {
        // post-stage-1 state, kept for the fix-ups; every chain starts pass 1 from a private copy of it
        double* orig = (double*)d_orig.p;
        B3D_CUDA(cudaMemcpyAsync(orig, poses_d, 8 * n * NPF, cudaMemcpyDeviceToDevice, st));
        B3D_CUDA(cudaMemcpyAsync(orig + n * NPF, R0_d, 8 * n * 9, cudaMemcpyDeviceToDevice, st));
        B3D_CUDA(cudaMemcpyAsync(orig + n * (NPF + 9), W_d, 8 * n * NS, cudaMemcpyDeviceToDevice, st));
        for (int c = 1; c < NC; ++c) {
            B3D_CUDA(cudaMemcpyAsync(kps_d + (size_t)c * n * MAXKP, kps_d, sizeof(b3d_kp) * n * MAXKP, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(nkp_d + (size_t)c * n, nkp_d, 4 * n, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(poses_d + (size_t)c * n * NPF, poses_d, 8 * n * NPF, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(R0_d + (size_t)c * n * 9, R0_d, 8 * n * 9, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(W_d + (size_t)c * n * NS, W_d, 8 * n * NS, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(r_d + (size_t)c * n * NJ * 9, r_d, 8 * n * NJ * 9, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(p_d + (size_t)c * n * 3, p_d, 8 * n * 3, cudaMemcpyDeviceToDevice, st));
        }
    }
    if (g_seq_waves.load() == 0 && waves > 0 && so->wsize < CH_MIN_FRAMES) {
        // every chain's windows back to back in ONE launch (cluster = chain), the write-back between two windows done by the kernel
        DevBuf d_order;
        if (int rc = d_order.alloc(sizeof(int32_t) * order.size())) return rc;
        B3D_CUDA(cudaMemcpyAsync(d_order.p, order.data(), sizeof(int32_t) * order.size(), cudaMemcpyHostToDevice, st));
        if (int rc = win_launch_chains(fc_d + 1, probs_d, NP, (const int32_t*)d_order.p, NC, waves, kps_d, nkp_d, R0_d, poses_d, W_d, r_d, p_d, us_d, ni_d, te_d,
                                       (double*)d_ws.p, st)) return rc;
        B3D_CUDA(cudaStreamSynchronize(st));          // d_order is released at the end of this scope
    } else {
        // wave i = the i-th window of every chain in one launch, the write-backs as separate launches (debug hook b3d_debug_seq_waves)
        for (int i = 0; i < waves; ++i) {
            const int p0 = wave_off[i], np = wave_off[i + 1] - p0;
            if (int rc = win_launch(fc_d + 1, probs_d + p0, np, kps_d, nkp_d, R0_d, poses_d, W_d, us_d + p0, nullptr, 0, ni_d + p0, te_d + p0,
                                    (double*)d_ws.p, st)) return rc;
            for (int j = 0; j < np; ++j)
                if (int rc = writeback_problem(probs[p0 + j], poses_d, R0_d, r_d, p_d)) return rc;
        }
    }
    if (sharded) {
        // ownership: chain (q, g) owns frames [start(k0), start(k1 of the same chain)) of its sequence (last shard: to the end)
        auto own_range = [&](int c, int& f0, int& f1) {
            const Chain& ch = chains[c];
            f0 = seq_first[ch.q] + starts[ch.q][ch.k0];
            f1 = (ch.k1 < (int)starts[ch.q].size()) ? seq_first[ch.q] + starts[ch.q][ch.k1] : seq_first[ch.q] + seq_len[ch.q];
        };
        auto copy_frames = [&](const double* spo, const double* sr0, const double* sw, const double* sr, const double* sp, int f0, int f1) -> int {
            if (f1 <= f0) return B3D_OK;
            const size_t a = (size_t)f0, cnt = (size_t)(f1 - f0);
            B3D_CUDA(cudaMemcpyAsync(poses_d + a * NPF, spo + a * NPF, 8 * cnt * NPF, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(R0_d + a * 9, sr0 + a * 9, 8 * cnt * 9, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(W_d + a * NS, sw + a * NS, 8 * cnt * NS, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(r_d + a * NJ * 9, sr + a * NJ * 9, 8 * cnt * NJ * 9, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(p_d + a * 3, sp + a * 3, 8 * cnt * 3, cudaMemcpyDeviceToDevice, st));
            return B3D_OK;
        };
        // pass 2 inputs: post-stage-1 state with the halo (overlap frames refined by the left neighbour) patched in.  The fix-up
        // of chain c runs in a scratch state f_* addressed through the same `first` row as chain c's private copy, so stage the
        // scratch INTO a separate buffer laid out like the private copies of the fix chains only.
        const double* orig = (const double*)d_orig.p;
        const size_t NFX = fix_chain.size();
        double* f_poses = (double*)d_fix.p; double* f_R0 = f_poses + NFX * n * NPF; double* f_W = f_R0 + NFX * n * 9;
        double* f_r = f_W + NFX * n * NS; double* f_p = f_r + NFX * n * NJ * 9;
        std::vector<WinProblem> fprobs(NFX);
        for (size_t j = 0; j < NFX; ++j) {
            const int c = fix_chain[j];
            const Chain& ch = chains[c];
            const int s = seq_first[ch.q] + starts[ch.q][ch.k0], e = std::min(s + so->wsize, seq_first[ch.q] + (int)seq_len[ch.q]), h = std::min(s + so->overlap, e);
            B3D_CUDA(cudaMemcpyAsync(f_poses + j * n * NPF, orig, 8 * n * NPF, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(f_R0 + j * n * 9, orig + n * NPF, 8 * n * 9, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemcpyAsync(f_W + j * n * NS, orig + n * (NPF + 9), 8 * n * NS, cudaMemcpyDeviceToDevice, st));
            B3D_CUDA(cudaMemsetAsync(f_r + j * n * NJ * 9, 0, 8 * n * NJ * 9, st));
            B3D_CUDA(cudaMemsetAsync(f_p + j * n * 3, 0, 8 * n * 3, st));
            if (h > s) {   // halo from chain c - 1 (the left neighbour of the same sequence)
                B3D_CUDA(cudaMemcpyAsync(f_poses + (j * n + s) * NPF, poses_d + ((size_t)(c - 1) * n + s) * NPF, 8 * (size_t)(h - s) * NPF, cudaMemcpyDeviceToDevice, st));
                B3D_CUDA(cudaMemcpyAsync(f_R0 + (j * n + s) * 9, R0_d + ((size_t)(c - 1) * n + s) * 9, 8 * (size_t)(h - s) * 9, cudaMemcpyDeviceToDevice, st));
            }
            fprobs[j] = {(int32_t)(j * n + s), e - s, probs[fix_off + j].ws_offset};
        }
        if (NFX) {
            B3D_CUDA(cudaMemcpyAsync(d_probs.p + sizeof(WinProblem) * (size_t)fix_off, fprobs.data(), sizeof(WinProblem) * NFX, cudaMemcpyHostToDevice, st));
            B3D_CUDA(cudaStreamSynchronize(st));
            // the fix chains read the keypoints through the same row index: kps copies are identical per copy, copy j < NC exists
            if (int rc = win_launch(fc_d + 1, probs_d + fix_off, (int)NFX, kps_d, nkp_d, f_R0, f_poses, f_W, us_d + fix_off, nullptr, 0,
                                    ni_d + fix_off, te_d + fix_off, (double*)d_ws.p, st)) return rc;
            for (size_t j = 0; j < NFX; ++j)
                if (int rc = writeback_problem(fprobs[j], f_poses, f_R0, f_r, f_p)) return rc;
        }
        // assemble the result in copy 0's arrays: owned frames from pass 1, then [s, keep_to) of every fix-up
        for (int c = 1; c < NC; ++c) {
            int f0, f1; own_range(c, f0, f1);
            const size_t b = (size_t)c * n;
            if (int rc = copy_frames(poses_d + b * NPF, R0_d + b * 9, W_d + b * NS, r_d + b * NJ * 9, p_d + b * 3, f0, f1)) return rc;
        }
        for (size_t j = 0; j < NFX; ++j) {
            const int c = fix_chain[j];
            const Chain& ch = chains[c];
            const int s = seq_first[ch.q] + starts[ch.q][ch.k0], e = std::min(s + so->wsize, seq_first[ch.q] + (int)seq_len[ch.q]);
            const int keep_to = (ch.k1 - ch.k0 > 1) ? std::min(s + stride, e) : e;   // frames the shard's 2nd window revisits keep pass-1 values
            if (int rc = copy_frames(f_poses + j * n * NPF, f_R0 + j * n * 9, f_W + j * n * NS, f_r + j * n * NJ * 9, f_p + j * n * 3, s, keep_to)) return rc;
        }
    }
    B3D_CUDA(cudaStreamSynchronize(st));
    B3D_CUDA(cudaMemcpy(poses, poses_d, 8 * n * NPF, cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(R0, R0_d, 8 * n * 9, cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(W, W_d, 8 * n * NS, cudaMemcpyDeviceToHost));
    if (r) B3D_CUDA(cudaMemcpy(r, r_d, 8 * n * NJ * 9, cudaMemcpyDeviceToHost));
    if (p) B3D_CUDA(cudaMemcpy(p, p_d, 8 * n * 3, cudaMemcpyDeviceToHost));
    std::vector<uint8_t> us(NP);
    std::vector<int32_t> ni(NP), te(NP);
    B3D_CUDA(cudaMemcpy(us.data(), us_d, NP, cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(ni.data(), ni_d, 4 * (size_t)NP, cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(te.data(), te_d, 4 * (size_t)NP, cudaMemcpyDeviceToHost));
    int nfail = 0;
    for (int i = 0; i < NP; ++i) nfail += us[i] ? 0 : 1;
    if (nfail_out) *nfail_out = nfail;
    // per-solve records in the oracle's order for a single sequence: [stage 1, pass-1 windows in index order, fix-ups in shard order]
    if (solve_iters || solve_term) {
        std::vector<std::pair<long long, int>> order;     // (sort key, problem)
        for (int q = 0; q < nseq; ++q) order.push_back({(long long)q, q});
        for (int i = 0; i < waves; ++i)
            for (size_t j = 0; j < wave_chain[i].size(); ++j) {
                const Chain& ch = chains[wave_chain[i][j]];
                order.push_back({(long long)nseq + (long long)ch.q * 1000000LL + ch.k0 + i, wave_off[i] + (int)j});
            }
        std::sort(order.begin() + nseq, order.end());
        for (size_t j = 0; j < fix_chain.size(); ++j) order.push_back({0, fix_off + (int)j});
        for (size_t i = 0; i < order.size(); ++i) {
            if (solve_iters) solve_iters[i] = ni[order[i].second];
            if (solve_term) solve_term[i] = te[order[i].second];
        }
    }
    return B3D_OK;
}

int b3d_fit_sequence(const b3d_model* m, const b3d_fit_opts* base, const b3d_seq_opts* so, int32_t N, const b3d_kp* kps, const int32_t* nkp,
                     double* poses, double* R0, double* W, double* r, double* p, int32_t* solve_iters, int32_t* solve_term, int32_t* nfail) {
    const int32_t first = 0, len = N;
    return b3d_fit_sequences(m, base, so, 1, &first, &len, N, kps, nkp, poses, R0, W, r, p, solve_iters, solve_term, nfail);
}

}  // extern "C"
