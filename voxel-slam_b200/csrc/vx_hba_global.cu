// vx_hba_global.cu — the bottom-up hierarchical global BA of VOXEL_SLAM::thd_globalmapping (voxelslam.cpp:3018-3141) as one
// call: window schedule, bottom-level HBA_add_edge of every window (batched, block-distributed over the ranks of a
// communicator), one submap per window (merged cloud, :2954-2985), top-level HBA_add_edge over the connected submaps
// (voxel-sharded across the ranks: exchange by root-voxel owner + all-reduce of every evaluation).
#include "vx_hba.cuh"
#include "vx_voxelize.cuh"
#include <chrono>
#include <algorithm>
#include <memory>
#include <map>

using namespace vx;

// ---- the schedule: a literal walk of the while-loop at voxelslam.cpp:3037-3139 over stored sessions -------------------------
extern "C" int vx_hba_schedule(const int32_t *session_sizes, int n_sessions, int wdsize, int mgsize, int32_t *win_session, int32_t *win_start,
                               int32_t *win_size, int cap, int32_t *n_windows) {
  VX_ARG(session_sizes && n_sessions >= 1 && wdsize >= 2 && mgsize >= 1 && mgsize <= wdsize && n_windows);
  int nw = 0;
  auto emit = [&](int s, int first, int count) -> int {
    if (count <= 0) return VX_OK;
    if (nw < cap && win_session && win_start && win_size) { win_session[nw] = s; win_start[nw] = first; win_size[nw] = count; }
    nw++;
    return VX_OK;
  };
  for (int s = 0; s < n_sessions; s++) {
    const int n = session_sizes[s];
    VX_ARG(n >= 0);
    int first = 0, count = 0;            // localID = [first, first + count)
    for (int buf_base = 0; buf_base < n;) {
      if (count == 0) first = buf_base;
      count++; buf_base++;               // localID.push_back(buf_base); buf_base++   (:3076-3078)
      if (count < wdsize) continue;      // :3079-3083
      emit(s, first, count);             // HBA_add_edge on the full window (:3096)
      if (buf_base >= n && s + 1 < n_sessions) { count = 0; break; }   // session used up and another follows: switch, localID.clear() (:3127-3133)
      first += mgsize; count -= mgsize;  // localID.pop_front() x mgsize (:3135-3138)
    }
    // session exhausted: the next session (:3127-3133) or the global pass (:3054-3058) takes the queue as it stands
    emit(s, first, count);
  }
  *n_windows = nw;
  if (nw > cap && win_session) { set_error("vx_hba_schedule: %d windows, capacity %d", nw, cap); return VX_ERR_ARG; }
  return VX_OK;
}

namespace {
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
// Windows of rank `rank`: a contiguous run, balanced by COST.  Full windows of a rank go through one batched build + one batched
// LM; every window of another size is a batch of its own, and a batch has a fixed cost (the host round trips of one build and
// its LM iterations: ~1.2 ms) worth about SHORT_WEIGHT full windows at the per-window rate of a many-window batch (0.06-0.07 ms).
// With plain counts the rank that holds a session's short tail window finished its bottom level 1.2 ms after the others and the
// whole job waited for it in the edge gather.
constexpr int SHORT_WEIGHT = 16;
void window_range(const int32_t *wsize, int n, int wdsize, int rank, int world, int &lo, int &hi) {
  int64_t total = 0;
  for (int w = 0; w < n; w++) total += wsize[w] == wdsize ? 1 : SHORT_WEIGHT;
  // window w goes to the rank whose share [r total / world, (r + 1) total / world) holds the midpoint of its weight interval
  auto owner = [&](int64_t before, int wt) { const int64_t mid2 = 2 * before + wt; return (int)std::min<int64_t>(world - 1, mid2 * world / (2 * total)); };
  lo = n; hi = n;
  int64_t acc = 0;
  bool open = false;
  for (int w = 0; w < n; w++) {
    const int wt = wsize[w] == wdsize ? 1 : SHORT_WEIGHT;
    const int r = owner(acc, wt);
    if (r == rank && !open) { lo = w; open = true; }
    if (r > rank && open) { hi = w; open = false; break; }
    if (r > rank && !open) { lo = hi = w; break; }
    acc += wt;
  }
  if (open) hi = n;
}
}  // namespace

extern "C" int vx_hba_window_range(const int32_t *win_size, int n_windows, int wdsize, int rank, int n_ranks, int32_t *lo, int32_t *hi) {
  VX_ARG(win_size && lo && hi && n_windows >= 0 && n_ranks >= 1 && rank >= 0 && rank < n_ranks);
  int a = 0, b = 0;
  window_range(win_size, n_windows, wdsize, rank, n_ranks, a, b);
  *lo = a; *hi = b;
  return VX_OK;
}

extern "C" int vx_hba_global(vx_ctx *ctx, vx_comm *comm, const float *xyz, const int64_t *cloud_off, int n_kf, const int32_t *kf_session,
                             const int32_t *kf_id, const vx_pose *kf_x0, const vx_pose *kf_x, const int32_t *cnct_map, int n_cnct,
                             const vx_hba_global_params *prm, vx_pgo_edge *edges1, int64_t cap1, int64_t *n1, vx_pgo_edge *edges2, int64_t cap2,
                             int64_t *n2, vx_pose *top_poses, int32_t *top_kf, int32_t *n_top, double *stats) {
  VX_ARG(ctx && xyz && cloud_off && n_kf >= 1 && kf_session && kf_x0 && prm && n1 && n2 && (comm == nullptr || comm->ctx == ctx));
  VX_ARG(prm->wdsize >= 2 && prm->mgsize >= 1 && prm->mgsize <= prm->wdsize && prm->bottom_max_iter >= 1 && prm->total_max_iter >= 0);
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  const int rank = comm ? comm->rank : 0, world = comm ? comm->size : 1;
  if (stats) for (int k = 0; k < 16; k++) stats[k] = 0.0;
  const double t_begin = now_ms();

  // ---- schedule ------------------------------------------------------------------------------------------------------------------
  std::vector<int32_t> sess_size, sess_first;
  for (int i = 0; i < n_kf; i++) {
    VX_ARG(kf_session[i] >= 0 && (i == 0 || kf_session[i] >= kf_session[i - 1]));
    while ((int)sess_size.size() <= kf_session[i]) { sess_size.push_back(0); sess_first.push_back(i); }
    sess_size[kf_session[i]]++;
  }
  const int n_sess = (int)sess_size.size();
  int32_t NWt = 0;
  VX_TRY(vx_hba_schedule(sess_size.data(), n_sess, prm->wdsize, prm->mgsize, nullptr, nullptr, nullptr, 0, &NWt));
  std::vector<int32_t> wsess(NWt), wstart(NWt), wsize(NWt);
  VX_TRY(vx_hba_schedule(sess_size.data(), n_sess, prm->wdsize, prm->mgsize, wsess.data(), wstart.data(), wsize.data(), NWt, &NWt));
  std::vector<int32_t> wkf(NWt);   // first keyframe (global index) of every window
  for (int w = 0; w < NWt; w++) { wkf[w] = sess_first[wsess[w]] + wstart[w]; VX_ARG(wsize[w] >= 2); }
  int lo = 0, hi = NWt;
  window_range(wsize.data(), NWt, prm->wdsize, rank, world, lo, hi);
  const int NWl = hi - lo;
  const int64_t ecap = (int64_t)prm->wdsize * (prm->wdsize - 1) / 2;

  // ---- this rank's keyframe clouds to the device (one contiguous range: windows are ordered by first keyframe) -------------------
  int kf_lo = n_kf, kf_hi = 0;
  for (int w = lo; w < hi; w++) { kf_lo = std::min(kf_lo, (int)wkf[w]); kf_hi = std::max(kf_hi, (int)wkf[w] + wsize[w]); }
  if (NWl == 0) { kf_lo = 0; kf_hi = 0; }
  const int n_lc = kf_hi - kf_lo;
  std::vector<int64_t> lcoff(n_lc + 1, 0);
  for (int c = 0; c <= n_lc; c++) lcoff[c] = cloud_off[kf_lo + c] - cloud_off[kf_lo];
  const int64_t NPl = lcoff[n_lc];
  DevBuf d_xyz, d_coff, d_sub, d_edges, d_tsub, d_tcoff;
  std::vector<DevBuf *> all = {&d_xyz, &d_coff, &d_sub, &d_edges, &d_tsub, &d_tcoff};
  std::unique_ptr<vx_pgo_edge[]> eall_own(new vx_pgo_edge[(size_t)NWt * ecap]);      // every slot that is read is written first (the multi-GPU gather zero-fills the rest itself)
  vx_pgo_edge *eall = eall_own.get();
  std::vector<int64_t> ecnt(NWt, 0);
  std::vector<int64_t> sub_off(NWl + 1, 0);
  int64_t vf_bottom[4] = {0, 0, 0, 0}, vf_top[4] = {0, 0, 0, 0}, top_points = 0;
  double t_upload = 0, t_bottom = 0, t_merge = 0, t_gather = 0, t_top = 0;
  int ntop = 0;

  auto body = [&]() -> int {
    double t0 = now_ms();
    // (Tried and removed: the full windows in 2-4 batches with the clouds of batch p + 1 crossing PCIe while batch p is optimised.
    //  Bit-identical edges once the fused Hessian kernel's item split was fixed, but slower: bench e2e 102.2 ms with one batch, 104.5
    //  with two, 107.2 with three — a batch's fixed cost (the host round trips of one build + three LM iterations) is larger than
    //  the share of the 8.6 ms upload it hides.)
    const float *d_kf = nullptr;                 // this rank's keyframe clouds on the device
    if (prm->xyz_on_device) d_kf = xyz + 3 * cloud_off[kf_lo];
    else {
      VX_TRY(d_xyz.reserve(12 * (size_t)std::max<int64_t>(NPl, 1)));
      if (NPl) VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz + 3 * cloud_off[kf_lo], 12 * (size_t)NPl, cudaMemcpyHostToDevice, st));
      d_kf = d_xyz.as<float>();
    }
    VX_TRY(d_coff.reserve(8 * (size_t)(n_lc + 1)));
    VX_CUDA(cudaMemcpyAsync(d_coff.p, lcoff.data(), 8 * (size_t)(n_lc + 1), cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaStreamSynchronize(st));
    t_upload = now_ms() - t0; t0 = now_ms();

    // ---- bottom level: windows grouped by size (all full windows in one batch; the short tails by their own size) -----------------
    std::vector<vx_pose> wposes;                 // optimized poses of this rank's windows, window-major (variable sizes)
    std::vector<int32_t> wf_off(NWl + 1, 0);
    for (int a = 0; a < NWl; a++) wf_off[a + 1] = wf_off[a] + wsize[lo + a];
    wposes.resize(wf_off[NWl]);
    std::vector<int32_t> fcloud(wf_off[NWl]), fmp(wf_off[NWl]);
    for (int a = 0; a < NWl; a++)
      for (int i = 0; i < wsize[lo + a]; i++) {
        const int kf = wkf[lo + a] + i;
        wposes[wf_off[a] + i] = kf_x0[kf];
        fcloud[wf_off[a] + i] = kf - kf_lo;
        fmp[wf_off[a] + i] = kf_session[kf];
      }
    std::map<int, std::vector<int>> by_size;
    for (int a = 0; a < NWl; a++) by_size[wsize[lo + a]].push_back(a);
    for (auto &kv : by_size) {
      const int W = kv.first;
      const std::vector<int> &grp = kv.second;
      const int ng = (int)grp.size();
      std::vector<vx_pose> ps((size_t)ng * W);
      std::vector<int32_t> ids((size_t)ng * W), mps((size_t)ng * W), fcl((size_t)ng * W);
      for (int b = 0; b < ng; b++)
        for (int i = 0; i < W; i++) {
          const int a = grp[b], kf = wkf[lo + a] + i;
          ps[(size_t)b * W + i] = wposes[wf_off[a] + i];
          ids[(size_t)b * W + i] = kf_id ? kf_id[kf] : kf - sess_first[kf_session[kf]];
          mps[(size_t)b * W + i] = kf_session[kf];
          fcl[(size_t)b * W + i] = kf - kf_lo;
        }
      const int64_t gcap = (int64_t)W * (W - 1) / 2;
      std::vector<vx_pgo_edge> ge((size_t)ng * gcap);
      std::vector<int64_t> gn(ng, 0);
      std::vector<int32_t> gst(ng, 0);
      HbaBuildFn build = [&](const std::vector<int> &ws, const vx_pose *p, const vx_gba_params &gp, vx_factor *fac) -> int {
        std::vector<int32_t> fc(ws.size() * (size_t)W);
        for (size_t b = 0; b < ws.size(); b++)
          for (int i = 0; i < W; i++) fc[b * W + i] = fcl[(size_t)ws[b] * W + i];
        return gba_build_impl(ctx, d_kf, d_coff.as<int64_t>(), lcoff.data(), n_lc, fc.data(), p, &gp, fac, nullptr, 0);
      };
      VX_TRY(hba_core(ctx, nullptr, build, ids.data(), mps.data(), ps.data(), W, ng, &prm->hba, prm->bottom_max_iter, prm->bottom_thread_num, ge.data(), gn.data(),
                      nullptr, nullptr, nullptr, gst.data(), vf_bottom, (prm->keep_factors && W == prm->wdsize) ? &ctx->kept[0] : nullptr));
      for (int b = 0; b < ng; b++) {
        const int a = grp[b], w = lo + a;
        if (gst[b] != VX_OK) { set_error("Too Less Voxel: bottom window %d (keyframes %d..%d)", w, wkf[w], wkf[w] + W - 1); return gst[b]; }
        for (int i = 0; i < W; i++) wposes[wf_off[a] + i] = ps[(size_t)b * W + i];
        ecnt[w] = gn[b];
        std::copy(ge.begin() + (size_t)b * gcap, ge.begin() + (size_t)b * gcap + gn[b], eall + (size_t)w * ecap);
      }
    }
    VX_CUDA(cudaStreamSynchronize(st));
    t_bottom = now_ms() - t0; t0 = now_ms();

    // ---- one submap per window: merged cloud in the frame of the window's first keyframe, down-sampled at voxel_size / 8 ---------
    if (NWl > 0) {
      int overflow = 0;
      VX_TRY(merge_clouds_batched(ctx, d_kf, lcoff.data(), n_lc, fcloud.data(), fmp.data(), wposes.data(), wf_off.data(), NWl,
                                  prm->hba.voxel_size / 8, d_sub, sub_off.data(), nullptr, &overflow));
    }
    d_xyz.release();
    VX_CUDA(cudaStreamSynchronize(st));
    t_merge = now_ms() - t0; t0 = now_ms();

    // ---- gba_edges1 of all ranks: every rank fills the slots of its windows in a zeroed array; an integer SUM all-reduce over
    //      the raw bytes is then an exact gather (x + 0 = x bit for bit)
    if (world > 1) {
      static_assert(sizeof(vx_pgo_edge) % 8 == 0, "edge records are whole int64 words");
      const size_t eb = sizeof(vx_pgo_edge) * (size_t)NWt * ecap, words = eb / 8 + NWt;
      VX_TRY(d_edges.reserve(8 * words));
      // the array is zeroed ON the device and only this rank's windows cross PCIe (their unused slots as zeros)
      for (int w = 0; w < NWt; w++) if (w < lo || w >= hi) ecnt[w] = 0;
      for (int w = lo; w < hi; w++) std::fill((char *)(&eall[(size_t)w * ecap] + ecnt[w]), (char *)(&eall[(size_t)w * ecap] + ecap), 0);
      VX_CUDA(cudaMemsetAsync(d_edges.p, 0, 8 * words, st));
      if (hi > lo) VX_CUDA(cudaMemcpyAsync(d_edges.as<char>() + sizeof(vx_pgo_edge) * (size_t)lo * ecap, eall + (size_t)lo * ecap, sizeof(vx_pgo_edge) * (size_t)(hi - lo) * ecap,
                                         cudaMemcpyHostToDevice, st));
      if (hi > lo) VX_CUDA(cudaMemcpyAsync(d_edges.as<char>() + eb + 8 * (size_t)lo, ecnt.data() + lo, 8 * (size_t)(hi - lo), cudaMemcpyHostToDevice, st));
      VX_TRY(comm_allreduce_i64(comm, d_edges.as<int64_t>(), (int64_t)words));
      VX_CUDA(cudaMemcpyAsync(eall, d_edges.p, eb, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaMemcpyAsync(ecnt.data(), d_edges.as<char>() + eb, 8 * (size_t)NWt, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
    }
    {
      int64_t tot = 0;
      for (int w = 0; w < NWt; w++) {
        if (edges1) {
          if (tot + ecnt[w] > cap1) { set_error("vx_hba_global: edges1 capacity %lld too small", (long long)cap1); return VX_ERR_ARG; }
          std::copy(eall + (size_t)w * ecap, eall + (size_t)w * ecap + ecnt[w], edges1 + tot);
        }
        tot += ecnt[w];
      }
      *n1 = tot;
    }
    t_gather = now_ms() - t0; t0 = now_ms();

    // ---- top level: the submaps of the connected sessions (voxelslam.cpp:2836-2857), xs = the scan pose of the submap's keyframe ----
    std::vector<int32_t> tw;   // window of every top-level frame
    {
      int last_mp = -1, isCnct = 0;
      for (int w = 0; w < NWt; w++) {
        const int mp = wsess[w];
        if (mp != last_mp) {
          isCnct = cnct_map == nullptr ? 1 : 0;
          for (int k = 0; k < n_cnct && cnct_map; k++) if (cnct_map[k] == mp) { isCnct = 1; break; }
          last_mp = mp;
        }
        if (isCnct) tw.push_back(w);
      }
    }
    ntop = (int)tw.size();
    *n2 = 0;
    if (n_top) *n_top = ntop;
    if (ntop >= 2 && prm->total_max_iter > 0) {
      std::vector<vx_pose> tp(ntop);
      std::vector<int32_t> tid(ntop), tmp(ntop);
      std::vector<int32_t> frame_of_window(NWt, -1);
      for (int g = 0; g < ntop; g++) {
        const int kf = wkf[tw[g]];
        tp[g] = kf_x ? kf_x[kf] : kf_x0[kf];
        tid[g] = kf_id ? kf_id[kf] : kf - sess_first[kf_session[kf]];
        tmp[g] = kf_session[kf];
        frame_of_window[tw[g]] = g;
      }
      // this rank's submap clouds that take part, as a compact cloud list
      std::vector<int64_t> toff(1, 0);
      std::vector<int32_t> tframe;
      bool compact = true;
      for (int a = 0; a < NWl; a++) if (frame_of_window[lo + a] < 0) compact = false;
      const float *d_tpts = d_sub.as<float>();
      if (compact) {
        for (int a = 0; a < NWl; a++) { toff.push_back(sub_off[a + 1]); tframe.push_back(frame_of_window[lo + a]); }
      } else {   // drop the submaps of unconnected sessions
        int64_t tot = 0;
        for (int a = 0; a < NWl; a++) if (frame_of_window[lo + a] >= 0) tot += sub_off[a + 1] - sub_off[a];
        VX_TRY(d_tsub.reserve(12 * (size_t)std::max<int64_t>(tot, 1)));
        int64_t at = 0;
        for (int a = 0; a < NWl; a++) {
          if (frame_of_window[lo + a] < 0) continue;
          const int64_t n = sub_off[a + 1] - sub_off[a];
          if (n) VX_CUDA(cudaMemcpyAsync(d_tsub.as<float>() + 3 * at, d_sub.as<float>() + 3 * sub_off[a], 12 * (size_t)n, cudaMemcpyDeviceToDevice, st));
          at += n; toff.push_back(at); tframe.push_back(frame_of_window[lo + a]);
        }
        d_tpts = d_tsub.as<float>();
      }
      const int n_tc = (int)tframe.size();
      top_points = toff.back();
      HbaBuildFn build;
      if (world > 1) {
        build = [&](const std::vector<int> &, const vx_pose *p, const vx_gba_params &gp, vx_factor *fac) -> int {
          return gba_build_sharded_impl(ctx, comm, d_tpts, toff.data(), n_tc, tframe.data(), p, &gp, fac, nullptr);
        };
      } else {
        VX_TRY(d_tcoff.reserve(8 * (size_t)(n_tc + 1)));
        VX_CUDA(cudaMemcpyAsync(d_tcoff.p, toff.data(), 8 * (size_t)(n_tc + 1), cudaMemcpyHostToDevice, st));
        std::vector<int32_t> fc(ntop);
        for (int c = 0; c < n_tc; c++) fc[tframe[c]] = c;
        build = [&, fc](const std::vector<int> &, const vx_pose *p, const vx_gba_params &gp, vx_factor *fac) -> int {
          return gba_build_impl(ctx, d_tpts, d_tcoff.as<int64_t>(), toff.data(), n_tc, fc.data(), p, &gp, fac, nullptr, 0);
        };
      }
      const int64_t tcap = (int64_t)ntop * (ntop - 1) / 2;
      PhaseTimer ptt("hba_global top", st);
      // the caller's edges2 takes the records directly when it can hold every pair; else an UNINITIALISED scratch (12.8 MB at 400 submaps:
      // value-initialising it cost 1-2.7 ms per job)
      std::unique_ptr<vx_pgo_edge[]> te_own;
      vx_pgo_edge *te = edges2;
      if (!edges2 || cap2 < tcap) { te_own.reset(new vx_pgo_edge[(size_t)tcap]); te = te_own.get(); }
      int64_t tn = 0;
      int32_t tst = VX_OK;
      VX_TRY(hba_core(ctx, comm, build, tid.data(), tmp.data(), tp.data(), ntop, 1, &prm->hba, prm->total_max_iter, prm->top_thread_num, te, &tn, nullptr, nullptr,
                      nullptr, &tst, vf_top, prm->keep_factors ? &ctx->kept[1] : nullptr));
      if (tst != VX_OK) { set_error("Too Less Voxel: top-level map over %d submaps", ntop); return tst; }
      if (edges2) {
        if (tn > cap2) { set_error("vx_hba_global: edges2 capacity %lld too small for %lld", (long long)cap2, (long long)tn); return VX_ERR_ARG; }
        if (te != edges2) std::copy(te, te + tn, edges2);
      }
      *n2 = tn;
      ptt.lap("hba_core + copy out");
      for (int g = 0; g < ntop; g++) { if (top_poses) top_poses[g] = tp[g]; if (top_kf) top_kf[g] = wkf[tw[g]]; }
    } else {
      for (int g = 0; g < ntop; g++) { const int kf = wkf[tw[g]]; if (top_poses) top_poses[g] = kf_x ? kf_x[kf] : kf_x0[kf]; if (top_kf) top_kf[g] = kf; }
    }
    VX_CUDA(cudaStreamSynchronize(st));
    t_top = now_ms() - t0;
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : all) b->release();
  if (stats) {
    stats[0] = NWt; stats[1] = ntop; stats[2] = (double)(vf_bottom[0] + vf_top[0]);
    stats[3] = t_upload; stats[4] = t_bottom; stats[5] = t_merge; stats[6] = t_gather; stats[7] = t_top; stats[8] = now_ms() - t_begin;
    stats[9] = (double)top_points; stats[10] = (double)vf_bottom[0]; stats[11] = (double)vf_top[0];
    stats[12] = (double)vf_bottom[1]; stats[13] = (double)vf_top[1]; stats[14] = (double)vf_top[2]; stats[15] = (double)vf_top[3];
  }
  return rc;
}

extern "C" int vx_hba_global_take_factor(vx_ctx *ctx, int which, vx_factor **out) {
  VX_ARG(ctx && out && (which == 0 || which == 1));
  *out = ctx->kept[which];
  ctx->kept[which] = nullptr;
  return VX_OK;
}

extern "C" int vx_factor_hessian_sharded_async(vx_ctx *ctx, vx_comm *comm, vx_factor *f) {
  VX_ARG(ctx && f && (comm == nullptr || comm->ctx == ctx));
  VX_TRY(vx_factor_hessian_async(ctx, f, 0, -1));
  return allreduce_evaluation(ctx, comm, f);
}
