// Workspace plan and the forward / backward launch sequences of the UNet (host only; every call enqueues kernels).
#include <string.h>

#include "common.cuh"
#include "host.cuh"
#include "unet.cuh"

namespace edb {

// ---------------------------------------------------------------------------------------------------------------
// workspace plan
// ---------------------------------------------------------------------------------------------------------------
struct Bump {
  char* base;
  size_t off = 0;
  template <class X>
  X* take(size_t n) {
    off = (off + 1023) & ~size_t(1023);
    X* p = reinterpret_cast<X*>(base + off);
    off += n * sizeof(X);
    return p;
  }
};

static size_t res_scratch(const ResLayer& l, int B) {
  const size_t m = static_cast<size_t>(B) * l.Hout * l.Wout;
  return (m * l.cout * 2 + m * (l.c0 + l.c1)) * 2 + 8192;   // da2, dh1, da1
}
static size_t st_scratch(const StLayer& l, int B) {
  const size_t m = static_cast<size_t>(B) * l.H * l.W;
  return m * l.C * 27 * 2 + static_cast<size_t>(B) * 32 * 2 * l.C * 4 + 65536;   // + cross-attention dK|dV scratch (T <= 32)   // dt3 dl3 dt2 datt2 dq2 dl2 dt1 datt dl1 dt0 dn (11) + dqkv (3) + dgg (4) + dg (8)
}

int unet_plan(UNet* u, int B, char* ws, bool training) {
  if (u->planB == B && u->plan_ws == ws && u->plan_training == training && ws != nullptr) return EDB_OK;
  const encdiff_unet_config& c = u->cfg;
  Bump a;
  a.base = ws;
  const int mc = c.model_channels, S = c.image_size;
  const size_t M0 = static_cast<size_t>(B) * S * S;
  u->temb = a.take<h16>(static_cast<size_t>(B) * mc);
  u->z1 = a.take<float>(static_cast<size_t>(B) * u->emb_ch);
  u->a1e = a.take<h16>(static_cast<size_t>(B) * u->emb_ch);
  u->z2 = a.take<float>(static_cast<size_t>(B) * u->emb_ch);
  u->a2e = a.take<h16>(static_cast<size_t>(B) * u->emb_ch);
  u->film = a.take<float>(static_cast<size_t>(B) * u->film_cols);
  u->cols = a.take<h16>(M0 * 64);
  u->aH = a.take<h16>(M0 * mc);
  u->stH = a.take<float>(static_cast<size_t>(B) * 64);
  u->yH = a.take<float>(M0 * c.out_channels);
  u->conv_stage.clear();
  u->ckv_dev = reinterpret_cast<CrossKvEntry*>(a.take<char>(64 * sizeof(CrossKvEntry)));
  u->ckv_dirty = true;
  if (training) {
    auto stage = [&](int param, long long Cp, long long Cq) {
      u->conv_stage.push_back(UNet::ConvStage{param, Cp, Cq, a.take<float>(static_cast<size_t>(9 * Cp * Cq))});
    };
    for (const ResLayer& l : u->res) {
      stage(l.conv1_w, l.cout, l.c0 + l.c1);
      stage(l.conv2_w, l.cout, l.cout);
    }
    stage(u->hc_w, c.out_channels, mc);
    u->fin_dev = reinterpret_cast<WgradFinalizeEntry*>(a.take<char>(u->conv_stage.size() * sizeof(WgradFinalizeEntry)));
    u->fin_dirty = true;
    u->wg_dev = reinterpret_cast<Wgrad2Params*>(a.take<char>(UNet::kWgCap * sizeof(Wgrad2Params) + 64));
    u->wg_dev = reinterpret_cast<Wgrad2Params*>((reinterpret_cast<uintptr_t>(u->wg_dev) + 63) & ~uintptr_t(63));
    u->wg_prefix_dev = reinterpret_cast<int*>(a.take<char>((UNet::kWgCap + 2) * sizeof(int)));
    u->cp_dev = reinterpret_cast<CrossProjEntry*>(a.take<char>(64 * sizeof(CrossProjEntry)));
    u->gs = a.take<float>(8);      // {scale, 1 / scale} + the two scratch words of grad_scale_kernel
    u->dfilm = a.take<h16>(static_cast<size_t>(B) * u->film_cols);
    u->dyH = a.take<h16>(M0 * 8);
  }
  // block-level tensors
  u->tn.clear();
  u->tn.reserve(1 + u->res.size() + u->st.size());
  auto new_tn = [&](int C, int H) {
    Tn t;
    t.C = C;
    t.H = t.W = H;
    t.p = a.take<h16>(static_cast<size_t>(B) * H * H * C);
    if (training) t.g = a.take<h16>(static_cast<size_t>(B) * H * H * C);
    u->tn.push_back(t);
    return &u->tn.back();
  };
  Tn* h = new_tn(mc, S);   // stem output
  u->hs_of_in_block.clear();
  std::vector<Tn*> hs;
  auto plan_layers = [&](std::vector<Layer>& ls, Tn* skip) {
    bool first = true;
    for (Layer& ly : ls) {
      if (ly.kind == 1) {
        ResLayer& l = u->res[ly.idx];
        const size_t mi = static_cast<size_t>(B) * l.Hin * l.Win, mo = static_cast<size_t>(B) * l.Hout * l.Wout;
        (void)mi;
        l.x0 = h;
        l.x1 = (first && skip) ? skip : nullptr;
        const int cin = l.c0 + l.c1;
        l.a1 = a.take<h16>(mo * cin);
        l.xs = (l.up || l.down) ? a.take<h16>(mo * cin) : nullptr;
        l.st1 = a.take<float>(static_cast<size_t>(B) * 64);
        l.h1 = a.take<h16>(mo * l.cout);
        l.st2 = a.take<float>(static_cast<size_t>(B) * 64);
        l.a2 = a.take<h16>(mo * l.cout);
        l.y = new_tn(l.cout, l.Hout);
        h = l.y;
      } else if (ly.kind == 2) {
        StLayer& l = u->st[ly.idx];
        const size_t m = static_cast<size_t>(B) * l.H * l.W;
        const int C = l.C;
        l.x = h;
        l.n = a.take<h16>(m * C);
        l.stn = a.take<float>(static_cast<size_t>(B) * 64);
        l.t0 = a.take<h16>(m * C);
        l.l1 = a.take<h16>(m * C);
        l.ls1 = a.take<float>(m * 2);
        l.qkv = a.take<h16>(m * 3 * C);
        l.att = a.take<h16>(m * C);
        l.lse = a.take<float>(m * l.heads);
        l.t1 = a.take<h16>(m * C);
        l.l2 = a.take<h16>(m * C);
        l.ls2 = a.take<float>(m * 2);
        l.q2 = a.take<h16>(m * C);
        l.att2 = a.take<h16>(m * C);
        l.t2 = a.take<h16>(m * C);
        l.l3 = a.take<h16>(m * C);
        l.ls3 = a.take<float>(m * 2);
        l.g = a.take<h16>(m * 8 * C);
        l.gg = a.take<h16>(m * 4 * C);
        l.t3 = a.take<h16>(m * C);
        l.kv = a.take<float>(static_cast<size_t>(B) * c.latent_unit * 2 * C);
        l.y = new_tn(C, l.H);
        h = l.y;
      }
      first = false;
    }
  };
  hs.push_back(h);
  for (size_t i = 1; i < u->in_blocks.size(); ++i) {
    plan_layers(u->in_blocks[i], nullptr);
    hs.push_back(h);
  }
  plan_layers(u->mid, nullptr);
  for (auto& blk : u->out_blocks) {
    Tn* skip = hs.back();
    hs.pop_back();
    plan_layers(blk, skip);
  }
  u->ws_bytes_persist = (a.off + 1023) & ~size_t(1023);
  size_t sc = static_cast<size_t>(B) * (u->emb_ch * 4) * 2 + M0 * (64 + mc) * 2 + 16384;   // emb MLP grads, dcols, daH
  if (training) {
    for (const ResLayer& l : u->res) sc = std::max(sc, res_scratch(l, B));
    for (const StLayer& l : u->st) sc = std::max(sc, st_scratch(l, B));
  } else {
    sc = 0;
  }
  sc = (sc + 1023) & ~size_t(1023);
  u->scratch_half = sc;
  u->ws_bytes_scratch = 2 * sc;      // two halves: the backward alternates per layer (side-stream readers, unet_bwd.inc)
  if (training && u->group_wgrad) {
    // deferred wgrads read the activation gradients after the whole chain has run: one slot per layer, never reused
    size_t tot = 3 * sc;             // head, stem, embedding MLP sections (each <= sc)
    for (const ResLayer& l : u->res) tot += (res_scratch(l, B) + 1023) & ~size_t(1023);
    for (const StLayer& l : u->st) tot += (st_scratch(l, B) + 1023) & ~size_t(1023);
    u->scratch_total = tot;
    u->ws_bytes_scratch = tot;
  }
  u->scratch = ws ? ws + u->ws_bytes_persist : nullptr;
  if (ws) {
    u->planB = B;
    u->plan_ws = ws;
    u->plan_training = training;
    u->act_maps.clear();
    u->out_maps.clear();
    u->grad_maps.clear();
  } else {
    u->planB = -1;
  }
  return EDB_OK;
}

size_t unet_workspace_bytes(UNet* u, int B, bool training) {
  // sizing runs the planner on a COPY: a live plan (layer pointers, tensor maps, device-table state) stays valid when a
  // caller merely asks for a size (bench.py does, between a trainer step and an autograd-path forward)
  UNet tmp = *u;
  unet_plan(&tmp, B, nullptr, training);
  return tmp.ws_bytes_persist + tmp.ws_bytes_scratch + 1024;
}

// ---------------------------------------------------------------------------------------------------------------
// launch helpers
// ---------------------------------------------------------------------------------------------------------------
struct Src {
  const h16* p;
  int C;
  int taps;
};

struct Exec {
  UNet* u;
  int B;
  cudaStream_t s;
  int err = 0;

  // ---- side stream (backward only): weight-gradient GEMMs overlap the input-gradient chain ------------------------
  // Layer k of the backward uses scratch half (k & 1).  A wgrad launched on the side stream reads activation
  // gradients of its layer's half, so the main stream waits for the side work of layer k-2 before layer k starts
  // writing that half again; join() folds the side stream back before the finalize kernel / the end of the call.
  int parity = 1;
  bool side_layer = false, side_any = false;
  bool done_valid[2] = {false, false};
  bool side_ready() {
    if (!u->side_on || u->prof) return false;      // profiling wants serialised, attributable launches
    if (!u->side) {
      if (cudaStreamCreateWithFlags(&u->side, cudaStreamNonBlocking) != cudaSuccess) return false;
      bool ok = true;
      for (cudaEvent_t* ev : {&u->ev_fork, &u->ev_join, &u->ev_side_done[0], &u->ev_side_done[1]})
        ok = ok && cudaEventCreateWithFlags(ev, cudaEventDisableTiming) == cudaSuccess;
      if (!ok) {
        u->side_on = false;
        return false;
      }
    }
    return true;
  }
  size_t bump = 0;       // group_wgrad mode: running offset into the never-reused scratch
  bool recording() const { return u->group_wgrad; }
  char* begin_layer(size_t bytes) {
    if (recording()) {
      char* p = u->scratch + bump;
      bump += (bytes + 1023) & ~size_t(1023);
      if (bump > u->scratch_total) invalid("internal: backward scratch overflow");
      return p;
    }
    return begin_layer();
  }
  char* begin_layer() {
    parity ^= 1;
    side_layer = false;
    if (done_valid[parity]) {
      cudaStreamWaitEvent(s, u->ev_side_done[parity], 0);
      done_valid[parity] = false;
    }
    return u->scratch + parity * u->scratch_half;
  }
  void end_layer() {
    if (side_layer) {
      cudaEventRecord(u->ev_side_done[parity], u->side);
      done_valid[parity] = true;
    }
  }
  cudaStream_t fork_side() {      // stream for work that only depends on what the main stream has enqueued so far
    if (!side_ready()) return s;
    cudaEventRecord(u->ev_fork, s);
    cudaStreamWaitEvent(u->side, u->ev_fork, 0);
    side_layer = side_any = true;
    return u->side;
  }
  // forward: first consumers of the FiLM table / projected context wait for the side stream's events (unet_forward)
  void wait_side_results(bool film, bool kv) {
    if (film && u->pend_film) {
      cudaStreamWaitEvent(s, u->ev_side_done[0], 0);
      u->pend_film = false;
    }
    if (kv && u->pend_kv) {
      cudaStreamWaitEvent(s, u->ev_side_done[1], 0);
      u->pend_kv = false;
    }
  }
  void join() {
    if (!side_any) return;
    cudaEventRecord(u->ev_join, u->side);
    cudaStreamWaitEvent(s, u->ev_join, 0);
    side_any = false;
    done_valid[0] = done_valid[1] = false;
  }

  double next_flops = 0;   // algorithmic FLOPs of the launch about to be reported through fail()
  double next_bytes = 0;   // algorithmic HBM bytes of it (operands read once + result written once)

  void mark(const char* what) {
    if (!u->prof) return;
    if (u->prof_used == u->prof_pool.size()) {
      cudaEvent_t ev;
      if (cudaEventCreate(&ev) != cudaSuccess) return;
      u->prof_pool.push_back(ev);
    }
    cudaEvent_t ev = u->prof_pool[u->prof_used++];
    cudaEventRecord(ev, s);
    u->prof_recs.push_back(ProfRec{what, next_flops, next_bytes, ev});
    next_flops = 0;
    next_bytes = 0;
  }

  // rc is a cudaError_t returned by a launcher
  bool fail(int rc, const char* what) {
    mark(what);
    if (rc != 0 && err == 0) {
      err = EDB_ERR_CUDA;
      set_last_error(std::string(what) + ": " + cudaGetErrorString(static_cast<cudaError_t>(rc)));
    }
    return err != 0;
  }
  void invalid(const char* what) {
    if (err == 0) {
      err = EDB_ERR_INVALID;
      set_last_error(what);
    }
  }

  const CUtensorMap* act_map(const h16* p, int C, int H, int W, const TileGeom& g) {
    auto key = std::make_tuple(static_cast<const void*>(p), C, H, W, B);
    auto it = u->act_maps.find(key);
    if (it != u->act_maps.end()) return &it->second;
    CUtensorMap m;
    int rc = make_act_tmap(&m, p, C, H, W, B, g);
    if (rc) {
      if (!err) err = rc;
      return nullptr;
    }
    return &u->act_maps.emplace(key, m).first->second;
  }
  const CUtensorMap* out_map(const void* p, long long M, int N, int ld) {
    auto key = std::make_tuple(p, M, N, ld);
    auto it = u->out_maps.find(key);
    if (it != u->out_maps.end()) return &it->second;
    CUtensorMap m;
    int rc = make_out_tmap(&m, p, M, N, ld);
    if (rc) {
      if (!err) err = rc;
      return nullptr;
    }
    return &u->out_maps.emplace(key, m).first->second;
  }
  const CUtensorMap* w_map(const h16* p, long long K, long long rows, int bn) {
    auto key = std::make_tuple(static_cast<const void*>(p), K, rows, bn);
    auto it = u->w_maps.find(key);
    if (it != u->w_maps.end()) return &it->second;
    CUtensorMap m;
    int rc = make_w_tmap(&m, p, K, rows, bn);
    if (rc) {
      if (!err) err = rc;
      return nullptr;
    }
    return &u->w_maps.emplace(key, m).first->second;
  }

  // out[M, N] = act(sum_s A_s * W^T + bias) + res ;  W = packed operand `pk` starting at row `row0`
  void gemm(std::initializer_list<Src> srcs, int H, int W, const Packed& pk, long long row0, int N, int tap_sign,
            const float* bias, const h16* res, int ldres, int act, bool fp32, void* out, int ld) {
    if (err) return;
    SgemmParams p;
    memset(&p, 0, sizeof(p));
    if (!make_geom(H, W, &p.g)) {
      invalid("gemm geometry: H*W must be a power of two, W <= 128");
      return;
    }
    int i = 0;
    long long K = 0;
    double a_bytes = 0;
    for (const Src& sc : srcs) {
      a_bytes += 2.0 * B * H * W * sc.C;
      const CUtensorMap* m = act_map(sc.p, sc.C, H, W, p.g);
      if (!m) return;
      p.amap[i] = *m;
      p.src_chunks[i] = (sc.C + 63) / 64;
      p.src_taps[i] = sc.taps;
      K += 64LL * p.src_chunks[i] * sc.taps;
      ++i;
    }
    if (K != pk.K) {
      invalid("internal: GEMM K mismatch with packed operand");
      return;
    }
    p.nsrc = i;
    const int sms = device_sm_count();
    const int m_tiles = (B * H * W + 127) / 128;
    const int bn2 = (!fp32 && ld == N && !u->force_old_gemm) ? sgemm2_pick_bn(m_tiles, N, sms, static_cast<int>(K / 64)) : 0;
    if (bn2) {
      Sgemm2Params q;
      memset(&q, 0, sizeof(q));
      for (int k = 0; k < 3; ++k) {
        q.amap[k] = p.amap[k];
        q.src_chunks[k] = p.src_chunks[k];
        q.src_taps[k] = p.src_taps[k];
      }
      q.nsrc = p.nsrc;
      q.g = p.g;
      const CUtensorMap* wm2 = w_map(u->pbf() + pk.off + row0 * pk.K, pk.K, pk.rows - row0, bn2);
      const CUtensorMap* om = out_map(out, static_cast<long long>(B) * H * W, N, ld);
      if (!wm2 || !om) return;
      q.bmap = *wm2;
      q.omap = *om;
      q.tap_sign = tap_sign;
      q.M = B * H * W;
      q.N = N;
      q.m_tiles = m_tiles;
      q.bias = bias;
      q.res = res;
      q.ldres = ldres;
      q.act = act;
      next_flops = 2.0 * q.M * static_cast<double>(N) * static_cast<double>(K);
      next_bytes = a_bytes + 2.0 * N * K + 2.0 * q.M * N * (res ? 2 : 1) + (bias ? 4.0 * N : 0.0);
      fail(launch_sgemm2(q, bn2, sms, s), "sgemm");
      return;
    }
    const int bn = sgemm_pick_bn(N);
    const CUtensorMap* wm = w_map(u->pbf() + pk.off + row0 * pk.K, pk.K, pk.rows - row0, bn);
    if (!wm) return;
    p.bmap = *wm;
    p.tap_sign = tap_sign;
    p.M = B * H * W;
    p.N = N;
    p.bias = bias;
    p.res = res;
    p.ldres = ldres;
    p.act = act;
    p.out_fp32 = fp32 ? 1 : 0;
    p.out0 = out;
    p.ld0 = ld;
    p.split_col = 1 << 30;
    next_flops = 2.0 * p.M * static_cast<double>(N) * static_cast<double>(K);
    next_bytes = a_bytes + 2.0 * N * K + (fp32 ? 4.0 : 2.0) * p.M * N + (res ? 2.0 * p.M * N : 0.0) + (bias ? 4.0 * N : 0.0);
    fail(launch_sgemm(p, bn, s), "sgemm.1tile");
  }

  void colsum(const h16* x, long long M, int ld, int col0, int N, float* d0, float* d1 = nullptr) {
    if (err || !d0) return;
    fail(launch_colsum(x, M, ld, col0, N, d0, d1, u->gs, s), "colsum");
  }

  const CUtensorMap* grad_map(float* base, int cols, int rows, bool three_d) {
    auto key = std::make_tuple(static_cast<const void*>(base), cols, rows, three_d ? 1 : 0);
    auto it = u->grad_maps.find(key);
    if (it != u->grad_maps.end()) return &it->second;
    CUtensorMap m;
    int rc = make_grad_tmap(&m, base, cols, rows, three_d);
    if (rc) {
      if (!err) err = rc;
      return nullptr;
    }
    return &u->grad_maps.emplace(key, m).first->second;
  }

  // grad[p, q(+q_off), tap] += sum_pix dy[pix, p_off + p] * x[pix, q];  bias0/bias1[p] += sum_pix dy[pix, p_off + p]
  void wgrad(const h16* dy, int Cdy, int p_off, int Cp, const h16* x, int Cx, int Cq, int H, int W, int taps,
             float* out, long long stride_p, long long stride_q, long long stride_tap, int q_off,
             float* bias0 = nullptr, float* bias1 = nullptr) {
    if (err || !out) return;
    // persistent kernel with TMA reduce-add epilogue and fused bias gradient (wgrad2.cu)
    float* dst = nullptr;
    bool three_d = false;
    int cols = 0;
    if (!u->force_old_gemm) {
      if (taps == 1 && stride_q == 1 && (stride_p % 4) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        dst = out;
        cols = static_cast<int>(stride_p);
      } else if (taps == 9 && stride_q == 9 && stride_tap == 1 && stride_p == 9LL * Cq && q_off == 0 && (Cq % 4) == 0) {
        for (const UNet::ConvStage& cs : u->conv_stage)
          if (u->gptr[cs.param] == out && cs.Cp == Cp && cs.Cq == Cq) dst = cs.ptr;
        three_d = true;
        cols = Cq;
      }
    }
    if (dst) {
      Wgrad2Params q;
      memset(&q, 0, sizeof(q));
      if (!make_geom(H, W, &q.g)) {
        invalid("wgrad geometry: H*W must be a power of two, W <= 128");
        return;
      }
      const CUtensorMap* pm = act_map(dy, Cdy, H, W, q.g);
      const CUtensorMap* qm = act_map(x, Cx, H, W, q.g);
      const CUtensorMap* om = grad_map(dst, cols, Cp, three_d);
      if (!pm || !qm || !om) return;
      q.pmap = *pm;
      q.qmap = *qm;
      q.omap = *om;
      q.taps = taps;
      q.m_tiles = (B * H * W + 127) / 128;
      q.Cp = Cp;
      q.Cq = Cq;
      q.p_off = p_off;
      q.q_off = q_off;
      q.out3d = three_d ? 1 : 0;
      q.bias0 = bias0;
      q.bias1 = bias1;
      q.gscale = u->gs;
      next_flops = 2.0 * B * H * W * static_cast<double>(Cp) * Cq * taps;
      if (recording()) {
        if (!u->fin_dirty) return;          // table already built for this plan / binding
        if (static_cast<int>(u->wg_host.size()) >= UNet::kWgCap) {
          invalid("internal: wgrad table capacity");
          return;
        }
        u->wg_prefix.push_back(u->wg_total);
        u->wg_total += wgrad2_prepare_group(&q, 16);
        u->wg_host.push_back(q);
        u->wg_dst.push_back(out);
        return;
      }
      fail(launch_wgrad2(q, device_sm_count(), fork_side()), "wgrad");
      return;
    }
    if (bias0) colsum(dy, static_cast<long long>(B) * H * W, Cdy, p_off, Cp, bias0, bias1);
    WgradParams p;
    memset(&p, 0, sizeof(p));
    if (!make_geom(H, W, &p.g)) {
      invalid("wgrad geometry: H*W must be a power of two, W <= 128");
      return;
    }
    const CUtensorMap* pm = act_map(dy, Cdy, H, W, p.g);
    const CUtensorMap* qm = act_map(x, Cx, H, W, p.g);
    if (!pm || !qm) return;
    p.pmap = *pm;
    p.qmap = *qm;
    p.taps = taps;
    p.m_tiles = (B * H * W + 127) / 128;
    p.Cp = Cp;
    p.Cq = Cq;
    p.p_off = p_off;
    p.out = out;
    p.stride_p = stride_p;
    p.stride_q = stride_q;
    p.stride_tap = stride_tap;
    p.q_off = q_off;
    p.gscale = u->gs;
    const int bn = Cq <= 64 ? 64 : 128;
    const long long ctas = static_cast<long long>((Cp + 127) / 128) * taps * ((Cq + bn - 1) / bn);
    // enough CTAs to fill the chip, but keep the number of fp32 atomics (weights x splits) bounded
    long long splits = (2 * 148 + ctas - 1) / ctas;
    const long long weights = static_cast<long long>(Cp) * Cq * taps;
    const long long max_by_atomics = std::max(1LL, (8LL << 20) / std::max(1LL, weights));
    splits = std::min(splits, max_by_atomics);
    splits = std::max(1LL, std::min<long long>(splits, p.m_tiles));
    next_flops = 2.0 * B * H * W * static_cast<double>(Cp) * Cq * taps;
    fail(launch_wgrad(p, bn, static_cast<int>(splits), s), "wgrad");
  }

  const float* P(int idx) const { return u->pptr[idx]; }
  float* G(int idx) const { return u->gptr.empty() ? nullptr : u->gptr[idx]; }
};

// ---------------------------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------------------------
static GnParams gn_res1(const ResLayer& l, Exec& e) {
  GnParams p;
  memset(&p, 0, sizeof(p));
  p.x0 = l.x0->p;
  p.C0 = l.c0;
  p.x1 = l.x1 ? l.x1->p : nullptr;
  p.C1 = l.c1;
  p.H = l.Hin;
  p.W = l.Win;
  p.gamma = e.P(l.gn1_w);
  p.beta = e.P(l.gn1_b);
  p.eps = 1e-5f;
  p.silu = 1;
  p.resample = l.down ? RESAMPLE_DOWN : (l.up ? RESAMPLE_UP : RESAMPLE_NONE);
  p.stats = l.st1;
  p.act = l.a1;
  p.xs = l.xs;
  return p;
}
static GnParams gn_res2(const ResLayer& l, Exec& e) {
  GnParams p;
  memset(&p, 0, sizeof(p));
  p.x0 = l.h1;
  p.C0 = l.cout;
  p.H = l.Hout;
  p.W = l.Wout;
  p.gamma = e.P(l.gn2_w);
  p.beta = e.P(l.gn2_b);
  p.film = e.u->film;
  p.ldfilm = e.u->film_cols;
  p.film_off = l.film_off;
  p.eps = 1e-5f;
  p.silu = 1;
  p.resample = RESAMPLE_NONE;
  p.stats = l.st2;
  p.act = l.a2;
  return p;
}

// ResBlock._forward (openaimodel_enc.py:255-275)
static void res_forward(ResLayer& l, Exec& e) {
  if (e.err) return;
  const int cin = l.c0 + l.c1;
  e.fail(launch_gn_fwd(gn_res1(l, e), e.B, e.s), "gn1");
  e.gemm({{l.a1, cin, 9}}, l.Hout, l.Wout, l.w1f, 0, l.cout, 1, e.P(l.conv1_b), nullptr, 0, ACT_NONE, false, l.h1, l.cout);
  e.wait_side_results(true, false);
  e.fail(launch_gn_fwd(gn_res2(l, e), e.B, e.s), "gn2");
  if (l.skipconv) {
    const float* bias = e.u->pf32() + l.bias2_off;
    if (l.x1)
      e.gemm({{l.a2, l.cout, 9}, {l.x0->p, l.c0, 1}, {l.x1->p, l.c1, 1}}, l.Hout, l.Wout, l.w2f, 0, l.cout, 1, bias,
             nullptr, 0, ACT_NONE, false, l.y->p, l.cout);
    else
      e.gemm({{l.a2, l.cout, 9}, {l.x0->p, l.c0, 1}}, l.Hout, l.Wout, l.w2f, 0, l.cout, 1, bias, nullptr, 0, ACT_NONE,
             false, l.y->p, l.cout);
  } else {
    const h16* skip = (l.up || l.down) ? l.xs : l.x0->p;
    e.gemm({{l.a2, l.cout, 9}}, l.Hout, l.Wout, l.w2f, 0, l.cout, 1, e.P(l.conv2_b), skip, l.cout, ACT_NONE, false,
           l.y->p, l.cout);
  }
}

// SpatialTransformer.forward + BasicTransformerBlock._forward (attention.py:211-215,250-261)
static void st_forward(StLayer& l, const float* ctx, Exec& e) {
  if (e.err) return;
  const int C = l.C, H = l.H, W = l.W, N = H * W;
  const int M = e.B * N;
  const encdiff_unet_config& c = e.u->cfg;
  GnParams p;
  memset(&p, 0, sizeof(p));
  p.x0 = l.x->p;
  p.C0 = C;
  p.H = H;
  p.W = W;
  p.gamma = e.P(l.norm_w);
  p.beta = e.P(l.norm_b);
  p.eps = 1e-6f;
  p.stats = l.stn;
  p.act = l.n;
  e.fail(launch_gn_fwd(p, e.B, e.s), "st.norm");
  e.gemm({{l.n, C, 1}}, H, W, l.pinf, 0, C, 1, e.P(l.pin_b), nullptr, 0, ACT_NONE, false, l.t0, C);
  // attn1 (self)
  e.fail(launch_ln_fwd(l.t0, e.P(l.ln1_w), e.P(l.ln1_b), 1e-5f, l.l1, l.ls1, M, C, e.s), "ln1");
  e.gemm({{l.l1, C, 1}}, H, W, l.qkvf, 0, 3 * C, 1, nullptr, nullptr, 0, ACT_NONE, false, l.qkv, 3 * C);
  e.fail(launch_self_attn_fwd(l.qkv, l.att, l.lse, e.B, N, l.heads, l.D, e.s), "self_attn");
  e.gemm({{l.att, C, 1}}, H, W, l.o1f, 0, C, 1, e.P(l.o1_b), l.t0, C, ACT_NONE, false, l.t1, C);
  // attn2 (cross, concept tokens)
  e.fail(launch_ln_fwd(l.t1, e.P(l.ln2_w), e.P(l.ln2_b), 1e-5f, l.l2, l.ls2, M, C, e.s), "ln2");
  e.gemm({{l.l2, C, 1}}, H, W, l.q2f, 0, C, 1, nullptr, nullptr, 0, ACT_NONE, false, l.q2, C);
  e.wait_side_results(false, true);
  e.fail(launch_cross_attn_fwd(l.q2, ctx, e.P(l.k2_w), e.P(l.v2_w), l.att2, e.B, N, l.heads, l.D, c.latent_unit,
                               c.context_dim, e.s, l.kv),
         "cross_attn");
  e.gemm({{l.att2, C, 1}}, H, W, l.o2f, 0, C, 1, e.P(l.o2_b), l.t1, C, ACT_NONE, false, l.t2, C);
  // GEGLU feed-forward
  e.fail(launch_ln_fwd(l.t2, e.P(l.ln3_w), e.P(l.ln3_b), 1e-5f, l.l3, l.ls3, M, C, e.s), "ln3");
  e.gemm({{l.l3, C, 1}}, H, W, l.ff1f, 0, 8 * C, 1, e.P(l.ff1_b), nullptr, 0, ACT_NONE, false, l.g, 8 * C);
  e.fail(launch_geglu_fwd(l.g, l.gg, M, 4 * C, e.s), "geglu");
  e.gemm({{l.gg, 4 * C, 1}}, H, W, l.ff2f, 0, C, 1, e.P(l.ff2_b), l.t2, C, ACT_NONE, false, l.t3, C);
  e.gemm({{l.t3, C, 1}}, H, W, l.poutf, 0, C, 1, e.P(l.pout_b), l.x->p, C, ACT_NONE, false, l.y->p, C);
}

int unet_forward(UNet* u, int B, const float* x, const float* noise, const float* sa, const float* sb, const float* t,
                 const float* ctx, float* out, float* xt, cudaStream_t s) {
  EDB_REQUIRE(u->bound, "forward before bind");
  EDB_REQUIRE(u->planB == B && u->plan_ws, "forward before plan");
  const encdiff_unet_config& c = u->cfg;
  Exec e{u, B, s};
  e.mark("(begin)");
  const int mc = c.model_channels, S = c.image_size, E = u->emb_ch;
  // The embedding chain (7 small launches, ~0.1 ms) does not depend on x: it runs on the side stream while the main stream
  // does the stem and the first GroupNorm / convolution; FiLM is first needed by the first ResBlock's second GroupNorm, the
  // projected context by the first SpatialTransformer.
  const bool forked = e.side_ready();
  if (forked) {
    cudaEventRecord(u->ev_fork, s);
    cudaStreamWaitEvent(u->side, u->ev_fork, 0);
    e.s = u->side;
  }
  // timestep embedding MLP + all 28 FiLM projections (openaimodel_enc.py:726-727, 218-224, 264)
  e.fail(launch_temb(t, u->temb, B, mc, e.s), "temb");
  e.gemm({{u->temb, mc, 1}}, 1, 1, u->te0f, 0, E, 1, e.P(u->te0_b), nullptr, 0, ACT_NONE, true, u->z1, E);
  e.fail(launch_silu_fwd(u->z1, u->a1e, static_cast<long long>(B) * E, e.s), "silu");
  e.gemm({{u->a1e, E, 1}}, 1, 1, u->te2f, 0, E, 1, e.P(u->te2_b), nullptr, 0, ACT_NONE, true, u->z2, E);
  e.fail(launch_silu_fwd(u->z2, u->a2e, static_cast<long long>(B) * E, e.s), "silu");
  e.gemm({{u->a2e, E, 1}}, 1, 1, u->filmf, 0, u->film_cols, 1, u->pf32() + u->film_bias_off, nullptr, 0, ACT_NONE, true,
         u->film, u->film_cols);
  if (forked) {
    cudaEventRecord(u->ev_side_done[0], e.s);
    u->pend_film = true;
  }
  // to_k / to_v of the context for all cross-attention layers at once (attention.py:174-176; one launch, not 16 x 8)
  EDB_REQUIRE(u->st.size() <= 64, "forward: more than 64 SpatialTransformer layers");
  if (!u->st.empty()) {
    if (u->ckv_dirty) {      // first forward after plan() / bind(): always an eager call (warm-up before any capture)
      u->ckv_host.clear();
      for (const StLayer& l : u->st) u->ckv_host.push_back(CrossKvEntry{e.P(l.k2_w), e.P(l.v2_w), l.kv, l.C, 0});
      EDB_CHECK_CUDA(cudaMemcpyAsync(u->ckv_dev, u->ckv_host.data(), u->ckv_host.size() * sizeof(CrossKvEntry),
                                     cudaMemcpyHostToDevice, s));
      EDB_CHECK_CUDA(cudaStreamSynchronize(s));
      u->ckv_dirty = false;
    }
    e.fail(launch_cross_kv_project(u->ckv_dev, static_cast<int>(u->st.size()), ctx, B, c.latent_unit, c.context_dim, e.s),
           "cross_attn.kv");
  }
  if (forked) {
    cudaEventRecord(u->ev_side_done[1], e.s);
    u->pend_kv = true;
    e.s = s;
  }
  // stem conv as one K=64 GEMM over the im2col'd 3-channel input (q_sample fused when noise != null)
  e.fail(launch_stem_im2col(x, noise, sa, sb, xt, u->cols, B, c.in_channels, S, S, s), "stem");
  e.gemm({{u->cols, 64, 1}}, S, S, u->stemf, 0, mc, 1, e.P(u->stem_b), nullptr, 0, ACT_NONE, false, u->tn[0].p, mc);
  auto run = [&](std::vector<Layer>& ls) {
    for (Layer& ly : ls) {
      if (ly.kind == 1) res_forward(u->res[ly.idx], e);
      else if (ly.kind == 2) st_forward(u->st[ly.idx], ctx, e);
    }
  };
  for (size_t i = 1; i < u->in_blocks.size(); ++i) run(u->in_blocks[i]);
  run(u->mid);
  for (auto& blk : u->out_blocks) run(blk);
  e.wait_side_results(true, true);      // (a network without ResBlocks / transformers: join before the call returns)
  // head: GN -> SiLU -> conv3x3 (openaimodel_enc.py:684-688,748)
  Tn& last = u->tn.back();
  GnParams p;
  memset(&p, 0, sizeof(p));
  p.x0 = last.p;
  p.C0 = last.C;
  p.H = p.W = S;
  p.gamma = e.P(u->hn_w);
  p.beta = e.P(u->hn_b);
  p.eps = 1e-5f;
  p.silu = 1;
  p.stats = u->stH;
  p.act = u->aH;
  e.fail(launch_gn_fwd(p, B, s), "head.gn");
  e.gemm({{u->aH, mc, 9}}, S, S, u->headf, 0, c.out_channels, 1, e.P(u->hc_b), nullptr, 0, ACT_NONE, true, u->yH,
         c.out_channels);
  e.fail(launch_nhwc_to_nchw(u->yH, out, B, c.out_channels, S, S, s), "head.out");
  return e.err ? EDB_ERR_CUDA : EDB_OK;
}

}  // namespace edb

#include "unet_bwd.inc"
#include "vq_enc.inc"
#include "enc4.inc"
