// Architecture walk, parameter inventory and packed-weight layout of the UNet (host only).
#include <string.h>

#include "host.cuh"
#include "unet.cuh"

namespace edb {

static long long ceil_to(long long v, long long m) { return (v + m - 1) / m * m; }

struct Builder {
  UNet* u;
  long long bf_off = 0;    // h16 arena cursor (elements)
  long long f32_off = 0;   // fp32 arena cursor (elements)

  int add_param(const std::string& name, std::initializer_list<long long> shape) {
    ParamInfo p;
    p.name = name;
    p.ndim = static_cast<int>(shape.size());
    p.numel = 1;
    int i = 0;
    for (long long s : shape) {
      p.shape[i++] = s;
      p.numel *= s;
    }
    for (; i < 4; ++i) p.shape[i] = 1;
    u->params.push_back(p);
    return static_cast<int>(u->params.size()) - 1;
  }

  Packed alloc(long long rows, long long K) {
    Packed pk;
    pk.rows = ceil_to(rows, 128);
    pk.K = K;
    pk.off = bf_off;
    bf_off += ceil_to(pk.rows * pk.K, 512);   // keep every operand 1 KB aligned
    return pk;
  }

  void entry(int kind, int src, int src2, long long dst_off, int rows, int cin, int taps, int cin_pad, int rows_pad,
             long long dst_ld, int row0, int col0) {
    PackEntry e;
    memset(&e, 0, sizeof(e));
    e.kind = kind;
    e.dst_off = dst_off;
    e.rows = rows;
    e.cin = cin;
    e.taps = taps;
    e.cin_pad = cin_pad;
    e.rows_pad = rows_pad;
    e.dst_ld = static_cast<int>(dst_ld);
    e.row0 = row0;
    e.col0 = col0;
    e.prefix = u->pack_total;
    u->pack_total += u->params[src].numel;
    e.blk0 = u->pack_blocks;
    u->pack_blocks += pack_entry_blocks(e);
    u->table.push_back(e);
    u->table_src.push_back(src);
    u->table_src2.push_back(src2);
  }

  // weight [rows, cin, taps] -> forward operand [rows_pad, taps*ceil64(cin)] placed at (row0, col0) of `dst`
  void pack_fwd(const Packed& dst, int src, int rows, int cin, int taps, int row0, int col0) {
    entry(0, src, -1, dst.off, rows, cin, taps, static_cast<int>(ceil_to(cin, 64)), 0, dst.K, row0, col0);
  }
  // transposed operand: dst[row0 + ci][col0 + tap*ceil64(rows) + r]
  void pack_t(const Packed& dst, int src, int rows, int cin, int taps, int row0, int col0) {
    entry(1, src, -1, dst.off, rows, cin, taps, 0, static_cast<int>(ceil_to(rows, 64)), dst.K, row0, col0);
  }
};

static void build_res(Builder& b, ResLayer& l, int emb_ch) {
  UNet* u = b.u;
  const std::string& p = l.prefix;
  const int cin = l.c0 + l.c1;
  l.gn1_w = b.add_param(p + ".in_layers.0.weight", {cin});
  l.gn1_b = b.add_param(p + ".in_layers.0.bias", {cin});
  l.conv1_w = b.add_param(p + ".in_layers.2.weight", {l.cout, cin, 3, 3});
  l.conv1_b = b.add_param(p + ".in_layers.2.bias", {l.cout});
  l.emb_w = b.add_param(p + ".emb_layers.1.weight", {2 * l.cout, emb_ch});
  l.emb_b = b.add_param(p + ".emb_layers.1.bias", {2 * l.cout});
  l.gn2_w = b.add_param(p + ".out_layers.0.weight", {l.cout});
  l.gn2_b = b.add_param(p + ".out_layers.0.bias", {l.cout});
  l.conv2_w = b.add_param(p + ".out_layers.3.weight", {l.cout, l.cout, 3, 3});
  l.conv2_b = b.add_param(p + ".out_layers.3.bias", {l.cout});
  l.skipconv = cin != l.cout;
  if (l.skipconv) {
    l.skip_w = b.add_param(p + ".skip_connection.weight", {l.cout, cin, 1, 1});
    l.skip_b = b.add_param(p + ".skip_connection.bias", {l.cout});
  }
  l.film_off = u->film_cols;
  u->film_cols += 2 * l.cout;
  const long long cin64 = ceil_to(cin, 64), cout64 = ceil_to(l.cout, 64);
  l.w1f = b.alloc(l.cout, 9 * cin64);
  b.pack_fwd(l.w1f, l.conv1_w, l.cout, cin, 9, 0, 0);
  l.w1t = b.alloc(cin, 9 * cout64);
  b.pack_t(l.w1t, l.conv1_w, l.cout, cin, 9, 0, 0);
  l.w2f = b.alloc(l.cout, 9 * cout64 + (l.skipconv ? cin64 : 0));
  b.pack_fwd(l.w2f, l.conv2_w, l.cout, l.cout, 9, 0, 0);
  l.w2t = b.alloc(l.cout, 9 * cout64);
  b.pack_t(l.w2t, l.conv2_w, l.cout, l.cout, 9, 0, 0);
  if (l.skipconv) {
    b.pack_fwd(l.w2f, l.skip_w, l.cout, cin, 1, 0, static_cast<int>(9 * cout64));
    l.wst = b.alloc(cin, cout64);
    b.pack_t(l.wst, l.skip_w, l.cout, cin, 1, 0, 0);
    l.bias2_off = b.f32_off;
    b.f32_off += ceil_to(l.cout, 64);
    b.entry(3, l.conv2_b, l.skip_b, l.bias2_off, l.cout, 1, 1, 0, 0, 0, 0, 0);
  }
}

static void build_st(Builder& b, StLayer& l, int ctx_dim) {
  const std::string& p = l.prefix;
  const int C = l.C;
  l.norm_w = b.add_param(p + ".norm.weight", {C});
  l.norm_b = b.add_param(p + ".norm.bias", {C});
  l.pin_w = b.add_param(p + ".proj_in.weight", {C, C, 1, 1});
  l.pin_b = b.add_param(p + ".proj_in.bias", {C});
  const std::string t = p + ".transformer_blocks.0";
  l.q1_w = b.add_param(t + ".attn1.to_q.weight", {C, C});
  l.k1_w = b.add_param(t + ".attn1.to_k.weight", {C, C});
  l.v1_w = b.add_param(t + ".attn1.to_v.weight", {C, C});
  l.o1_w = b.add_param(t + ".attn1.to_out.0.weight", {C, C});
  l.o1_b = b.add_param(t + ".attn1.to_out.0.bias", {C});
  l.ff1_w = b.add_param(t + ".ff.net.0.proj.weight", {8 * C, C});
  l.ff1_b = b.add_param(t + ".ff.net.0.proj.bias", {8 * C});
  l.ff2_w = b.add_param(t + ".ff.net.2.weight", {C, 4 * C});
  l.ff2_b = b.add_param(t + ".ff.net.2.bias", {C});
  l.q2_w = b.add_param(t + ".attn2.to_q.weight", {C, C});
  l.k2_w = b.add_param(t + ".attn2.to_k.weight", {C, ctx_dim});
  l.v2_w = b.add_param(t + ".attn2.to_v.weight", {C, ctx_dim});
  l.o2_w = b.add_param(t + ".attn2.to_out.0.weight", {C, C});
  l.o2_b = b.add_param(t + ".attn2.to_out.0.bias", {C});
  l.ln1_w = b.add_param(t + ".norm1.weight", {C});
  l.ln1_b = b.add_param(t + ".norm1.bias", {C});
  l.ln2_w = b.add_param(t + ".norm2.weight", {C});
  l.ln2_b = b.add_param(t + ".norm2.bias", {C});
  l.ln3_w = b.add_param(t + ".norm3.weight", {C});
  l.ln3_b = b.add_param(t + ".norm3.bias", {C});
  l.pout_w = b.add_param(p + ".proj_out.weight", {C, C, 1, 1});
  l.pout_b = b.add_param(p + ".proj_out.bias", {C});
  auto lin = [&](Packed& f, Packed& tt, int src, int rows, int cin) {
    f = b.alloc(rows, ceil_to(cin, 64));
    b.pack_fwd(f, src, rows, cin, 1, 0, 0);
    tt = b.alloc(cin, ceil_to(rows, 64));
    b.pack_t(tt, src, rows, cin, 1, 0, 0);
  };
  lin(l.pinf, l.pint, l.pin_w, C, C);
  l.qkvf = b.alloc(3 * C, C);
  l.qkvt = b.alloc(C, 3 * C);
  const int qkv_src[3] = {l.q1_w, l.k1_w, l.v1_w};
  for (int i = 0; i < 3; ++i) {
    b.pack_fwd(l.qkvf, qkv_src[i], C, C, 1, i * C, 0);
    b.pack_t(l.qkvt, qkv_src[i], C, C, 1, 0, i * C);
  }
  lin(l.o1f, l.o1t, l.o1_w, C, C);
  lin(l.q2f, l.q2t, l.q2_w, C, C);
  lin(l.o2f, l.o2t, l.o2_w, C, C);
  lin(l.ff1f, l.ff1t, l.ff1_w, 8 * C, C);
  lin(l.ff2f, l.ff2t, l.ff2_w, C, 4 * C);
  lin(l.poutf, l.poutt, l.pout_w, C, C);
}

int unet_build(UNet* u, const encdiff_unet_config* c) {
  u->cfg = *c;
  EDB_REQUIRE(c->model_channels > 0 && c->model_channels % 64 == 0, "model_channels must be a multiple of 64");
  EDB_REQUIRE(c->use_scale_shift_norm == 1, "only use_scale_shift_norm=True is implemented");
  EDB_REQUIRE(c->resblock_updown == 1, "only resblock_updown=True is implemented");
  EDB_REQUIRE(c->transformer_depth == 1, "only transformer_depth=1 is implemented");
  EDB_REQUIRE(c->in_channels * 9 <= 64, "in_channels*9 must fit one 64-wide K chunk");
  EDB_REQUIRE(c->out_channels <= 8, "out_channels must be <= 8");
  EDB_REQUIRE(c->n_channel_mult >= 1 && c->n_channel_mult <= 8 && c->n_attention_resolutions <= 8, "bad config");
  EDB_REQUIRE(c->num_heads > 0 && c->latent_unit >= 1 && c->latent_unit <= 32, "latent_unit must be in 1..32");
  Builder b;
  b.u = u;
  const int mc = c->model_channels;
  u->emb_ch = 4 * mc;
  u->te0_w = b.add_param("time_embed.0.weight", {u->emb_ch, mc});
  u->te0_b = b.add_param("time_embed.0.bias", {u->emb_ch});
  u->te2_w = b.add_param("time_embed.2.weight", {u->emb_ch, u->emb_ch});
  u->te2_b = b.add_param("time_embed.2.bias", {u->emb_ch});
  auto in_attn = [&](int ds) {
    for (int i = 0; i < c->n_attention_resolutions; ++i)
      if (c->attention_resolutions[i] == ds) return true;
    return false;
  };
  auto new_res = [&](const std::string& prefix, int c0, int c1, int cout, bool up, bool down, int H) {
    ResLayer l;
    l.prefix = prefix;
    l.c0 = c0;
    l.c1 = c1;
    l.cout = cout;
    l.up = up;
    l.down = down;
    l.Hin = l.Win = H;
    l.Hout = l.Wout = up ? 2 * H : (down ? H / 2 : H);
    u->res.push_back(l);
    return Layer{1, static_cast<int>(u->res.size()) - 1};
  };
  auto new_st = [&](const std::string& prefix, int C, int H) {
    StLayer l;
    l.prefix = prefix;
    l.C = C;
    l.heads = c->num_heads;
    l.D = C / c->num_heads;
    l.H = l.W = H;
    u->st.push_back(l);
    return Layer{2, static_cast<int>(u->st.size()) - 1};
  };
  // ---- structure (openaimodel_enc.py:516-682) -----------------------------------------------------------------
  std::vector<int> chans;
  int ch = mc, ds = 1, H = c->image_size;
  u->in_blocks.push_back({Layer{0, 0}});
  chans.push_back(ch);
  for (int level = 0; level < c->n_channel_mult; ++level) {
    const int mult = c->channel_mult[level];
    for (int r = 0; r < c->num_res_blocks; ++r) {
      const std::string pre = "input_blocks." + std::to_string(u->in_blocks.size());
      std::vector<Layer> blk;
      blk.push_back(new_res(pre + ".0", ch, 0, mult * mc, false, false, H));
      ch = mult * mc;
      if (in_attn(ds)) blk.push_back(new_st(pre + ".1", ch, H));
      u->in_blocks.push_back(blk);
      chans.push_back(ch);
    }
    if (level != c->n_channel_mult - 1) {
      const std::string pre = "input_blocks." + std::to_string(u->in_blocks.size());
      u->in_blocks.push_back({new_res(pre + ".0", ch, 0, ch, false, true, H)});
      chans.push_back(ch);
      ds *= 2;
      H /= 2;
    }
  }
  u->mid.push_back(new_res("middle_block.0", ch, 0, ch, false, false, H));
  u->mid.push_back(new_st("middle_block.1", ch, H));
  u->mid.push_back(new_res("middle_block.2", ch, 0, ch, false, false, H));
  for (int level = c->n_channel_mult - 1; level >= 0; --level) {
    const int mult = c->channel_mult[level];
    for (int i = 0; i <= c->num_res_blocks; ++i) {
      const int ich = chans.back();
      chans.pop_back();
      const std::string pre = "output_blocks." + std::to_string(u->out_blocks.size());
      std::vector<Layer> blk;
      blk.push_back(new_res(pre + ".0", ch, ich, mc * mult, false, false, H));
      ch = mc * mult;
      if (in_attn(ds)) blk.push_back(new_st(pre + ".1", ch, H));
      if (level && i == c->num_res_blocks) {
        blk.push_back(new_res(pre + "." + std::to_string(blk.size()), ch, 0, ch, true, false, H));
        ds /= 2;
        H *= 2;
      }
      u->out_blocks.push_back(blk);
    }
  }
  u->head_ch = ch;
  // ---- parameters in reference registration order + packed layout ---------------------------------------------
  u->stem_w = b.add_param("input_blocks.0.0.weight", {mc, c->in_channels, 3, 3});
  u->stem_b = b.add_param("input_blocks.0.0.bias", {mc});
  auto build_layers = [&](std::vector<Layer>& ls) -> int {
    for (Layer& l : ls) {
      if (l.kind == 1) {
        ResLayer& r = u->res[l.idx];
        if (r.c1 > 0 && (r.c0 % 64) != 0) return 1;
        build_res(b, r, u->emb_ch);
      } else if (l.kind == 2) {
        StLayer& s = u->st[l.idx];
        if (s.D != 8 && s.D != 16 && s.D != 32 && s.D != 64) return 2;
        build_st(b, s, c->context_dim);
      }
    }
    return 0;
  };
  for (size_t i = 1; i < u->in_blocks.size(); ++i)
    EDB_REQUIRE(build_layers(u->in_blocks[i]) == 0, "unsupported channel / head configuration");
  EDB_REQUIRE(build_layers(u->mid) == 0, "unsupported channel / head configuration");
  for (auto& blk : u->out_blocks) EDB_REQUIRE(build_layers(blk) == 0, "unsupported channel / head configuration");
  u->hn_w = b.add_param("out.0.weight", {ch});
  u->hn_b = b.add_param("out.0.bias", {ch});
  u->hc_w = b.add_param("out.2.weight", {c->out_channels, mc, 3, 3});
  u->hc_b = b.add_param("out.2.bias", {c->out_channels});
  EDB_REQUIRE(ch == mc, "head expects model_channels inputs (channel_mult[0] must be 1)");
  // stem: reference-order flat K (ci*9 + tap), one 64-wide chunk
  const int k_stem = c->in_channels * 9;
  u->stemf = b.alloc(mc, 64);
  b.entry(4, u->stem_w, -1, u->stemf.off, mc, c->in_channels, 9, 0, 0, 64, 0, 0);
  u->stemt = b.alloc(k_stem, ceil_to(mc, 64));
  b.entry(5, u->stem_w, -1, u->stemt.off, mc, c->in_channels, 9, 0, 0, u->stemt.K, 0, 0);
  // time embedding MLP
  auto lin = [&](Packed& f, Packed& tt, int src, int rows, int cin) {
    f = b.alloc(rows, ceil_to(cin, 64));
    b.pack_fwd(f, src, rows, cin, 1, 0, 0);
    tt = b.alloc(cin, ceil_to(rows, 64));
    b.pack_t(tt, src, rows, cin, 1, 0, 0);
  };
  lin(u->te0f, u->te0t, u->te0_w, u->emb_ch, mc);
  lin(u->te2f, u->te2t, u->te2_w, u->emb_ch, u->emb_ch);
  // all ResBlock emb_layers stacked into one [F, emb_ch] operand (K2 of SURVEY.md §2.3)
  u->filmf = b.alloc(u->film_cols, u->emb_ch);
  u->filmt = b.alloc(u->emb_ch, u->film_cols);
  u->film_bias_off = b.f32_off;
  b.f32_off += ceil_to(u->film_cols, 64);
  for (ResLayer& r : u->res) {
    b.pack_fwd(u->filmf, r.emb_w, 2 * r.cout, u->emb_ch, 1, r.film_off, 0);
    b.pack_t(u->filmt, r.emb_w, 2 * r.cout, u->emb_ch, 1, 0, r.film_off);
    b.entry(2, r.emb_b, -1, u->film_bias_off, 2 * r.cout, 1, 1, 0, 0, 0, 0, r.film_off);
  }
  // head conv
  u->headf = b.alloc(c->out_channels, 9 * ceil_to(mc, 64));
  b.pack_fwd(u->headf, u->hc_w, c->out_channels, mc, 9, 0, 0);
  u->headt = b.alloc(mc, 9 * 64);
  b.pack_t(u->headt, u->hc_w, c->out_channels, mc, 9, 0, 0);

  u->table_bytes = static_cast<size_t>(ceil_to(static_cast<long long>(u->table.size() * sizeof(PackEntry)), 1024));
  u->h16_bytes = static_cast<size_t>(ceil_to(b.bf_off * 2, 1024));
  u->f32_bytes = static_cast<size_t>(ceil_to(b.f32_off * 4, 1024));
  return EDB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// First-stage VQ-f4 encoder (ldm/modules/diffusionmodules/model.py:368-459 + quant_conv, autoencoder.py:49,313-316):
// parameter inventory in the reference's registration order and forward-layout operands only (frozen, inference).
// ---------------------------------------------------------------------------------------------------------------
int vqenc_build(UNet* u, const encdiff_vqenc_config* c) {
  u->model_kind = 1;
  u->vqcfg = *c;
  memset(&u->cfg, 0, sizeof(u->cfg));
  EDB_REQUIRE(c->ch >= 32 && c->ch % 32 == 0, "vqenc: ch must be a multiple of 32 (GroupNorm 32)");
  EDB_REQUIRE(c->n_ch_mult >= 1 && c->n_ch_mult <= 4 && c->num_res_blocks >= 1, "vqenc: bad ch_mult / num_res_blocks");
  EDB_REQUIRE(c->in_channels * 9 <= 64, "vqenc: in_channels*9 must fit one 64-wide K chunk");
  EDB_REQUIRE(c->z_channels >= 1 && c->z_channels <= 8 && c->embed_dim >= 1 && c->embed_dim <= 8, "vqenc: z_channels / embed_dim <= 8");
  int R = c->resolution;
  EDB_REQUIRE(R >= 8 && R <= 128 && (R & (R - 1)) == 0, "vqenc: resolution must be a power of two in 8..128");
  Builder b;
  b.u = u;
  auto conv_params = [&](const std::string& n, int co, int ci, int k, int& w, int& bi) {
    w = b.add_param(n + ".weight", {co, ci, k, k});
    bi = b.add_param(n + ".bias", {co});
  };
  auto norm_params = [&](const std::string& n, int ch, int& w, int& bi) {
    w = b.add_param(n + ".weight", {ch});
    bi = b.add_param(n + ".bias", {ch});
  };
  auto add_res = [&](const std::string& n, int ci, int co, int H) {
    VqLayer l;
    l.kind = 1;
    l.cin = ci;
    l.cout = co;
    l.H = H;
    norm_params(n + ".norm1", ci, l.n1w, l.n1b);
    conv_params(n + ".conv1", co, ci, 3, l.c1w, l.c1b);
    norm_params(n + ".norm2", co, l.n2w, l.n2b);
    conv_params(n + ".conv2", co, co, 3, l.c2w, l.c2b);
    if (ci != co) conv_params(n + ".nin_shortcut", co, ci, 1, l.sw, l.sb);
    const long long ci64 = ceil_to(ci, 64), co64 = ceil_to(co, 64);
    l.w1 = b.alloc(co, 9 * ci64);
    b.pack_fwd(l.w1, l.c1w, co, ci, 9, 0, 0);
    l.w2 = b.alloc(co, 9 * co64 + (l.sw >= 0 ? ci64 : 0));
    b.pack_fwd(l.w2, l.c2w, co, co, 9, 0, 0);
    if (l.sw >= 0) {
      b.pack_fwd(l.w2, l.sw, co, ci, 1, 0, static_cast<int>(9 * co64));
      l.bias_off = b.f32_off;
      b.f32_off += ceil_to(co, 64);
      b.entry(3, l.c2b, l.sb, l.bias_off, co, 1, 1, 0, 0, 0, 0, 0);
    }
    u->vq.push_back(l);
  };
  conv_params("encoder.conv_in", c->ch, c->in_channels, 3, u->vq_in_w, u->vq_in_b);
  int block_in = c->ch;
  for (int lvl = 0; lvl < c->n_ch_mult; ++lvl) {
    block_in = c->ch * (lvl == 0 ? 1 : c->ch_mult[lvl - 1]);
    const int block_out = c->ch * c->ch_mult[lvl];
    EDB_REQUIRE(block_out % 32 == 0, "vqenc: channel counts must be multiples of 32");
    for (int i = 0; i < c->num_res_blocks; ++i) {
      add_res("encoder.down." + std::to_string(lvl) + ".block." + std::to_string(i), block_in, block_out, R);
      block_in = block_out;
    }
    if (lvl != c->n_ch_mult - 1) {
      VqLayer l;
      l.kind = 2;
      l.cin = l.cout = block_in;
      l.H = R;
      conv_params("encoder.down." + std::to_string(lvl) + ".downsample.conv", block_in, block_in, 3, l.c1w, l.c1b);
      l.w1 = b.alloc(block_in, 9 * ceil_to(block_in, 64));
      b.pack_fwd(l.w1, l.c1w, block_in, block_in, 9, 0, 0);
      u->vq.push_back(l);
      R /= 2;
    }
  }
  EDB_REQUIRE(R >= 4 && R * R <= 256 && block_in <= 128 && block_in % 64 == 0,
              "vqenc: the mid attention kernel handles <= 256 tokens of 64 or 128 channels");
  add_res("encoder.mid.block_1", block_in, block_in, R);
  {
    VqLayer l;
    l.kind = 3;
    l.cin = l.cout = block_in;
    l.H = R;
    const std::string n = "encoder.mid.attn_1";
    norm_params(n + ".norm", block_in, l.n1w, l.n1b);
    conv_params(n + ".q", block_in, block_in, 1, l.qw, l.qb);
    conv_params(n + ".k", block_in, block_in, 1, l.kw, l.kb);
    conv_params(n + ".v", block_in, block_in, 1, l.vw, l.vb);
    conv_params(n + ".proj_out", block_in, block_in, 1, l.pw, l.pb);
    const int C = block_in;
    l.w1 = b.alloc(3 * C, ceil_to(C, 64));
    const int src[3] = {l.qw, l.kw, l.vw}, bsrc[3] = {l.qb, l.kb, l.vb};
    l.bias_off = b.f32_off;
    b.f32_off += ceil_to(3 * C, 64);
    for (int i = 0; i < 3; ++i) {
      b.pack_fwd(l.w1, src[i], C, C, 1, i * C, 0);
      b.entry(2, bsrc[i], -1, l.bias_off, C, 1, 1, 0, 0, 0, 0, i * C);
    }
    l.w2 = b.alloc(C, ceil_to(C, 64));
    b.pack_fwd(l.w2, l.pw, C, C, 1, 0, 0);
    u->vq.push_back(l);
  }
  add_res("encoder.mid.block_2", block_in, block_in, R);
  norm_params("encoder.norm_out", block_in, u->vq_no_w, u->vq_no_b);
  conv_params("encoder.conv_out", c->z_channels, block_in, 3, u->vq_co_w, u->vq_co_b);
  conv_params("quant_conv", c->embed_dim, c->z_channels, 1, u->vq_qc_w, u->vq_qc_b);
  u->vq_out_res = R;
  u->vq_out_ch = block_in;
  u->vq_order.clear();
  for (size_t i = 0; i < u->vq.size(); ++i) u->vq_order.push_back(static_cast<int>(i));
  // conv_in: reference-order flat K (ci*9 + tap) in one 64-wide chunk, like the UNet stem
  u->stemf = b.alloc(c->ch, 64);
  b.entry(4, u->vq_in_w, -1, u->stemf.off, c->ch, c->in_channels, 9, 0, 0, 64, 0, 0);
  u->headf = b.alloc(c->z_channels, 9 * ceil_to(block_in, 64));
  b.pack_fwd(u->headf, u->vq_co_w, c->z_channels, block_in, 9, 0, 0);
  u->table_bytes = static_cast<size_t>(ceil_to(static_cast<long long>(u->table.size() * sizeof(PackEntry)), 1024));
  u->h16_bytes = static_cast<size_t>(ceil_to(b.bf_off * 2, 1024));
  u->f32_bytes = static_cast<size_t>(ceil_to(b.f32_off * 4, 1024));
  return EDB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// First-stage decoder (model.py:462-568) + quantize.embedding + post_quant_conv (autoencoder.py:44-53).
// Parameters in the reference's registration order (conv_in, mid, up.0 .. up.L-1, norm_out, conv_out, then
// quantize.embedding.weight, post_quant_conv.*); execution order is mid, then the levels from the deepest one.
// ---------------------------------------------------------------------------------------------------------------
int vqdec_build(UNet* u, const encdiff_vqdec_config* cd) {
  const encdiff_vqenc_config* c = &cd->dd;
  u->model_kind = 2;
  u->vqcfg = *c;
  memset(&u->cfg, 0, sizeof(u->cfg));
  EDB_REQUIRE(c->ch >= 32 && c->ch % 32 == 0, "vqdec: ch must be a multiple of 32 (GroupNorm 32)");
  EDB_REQUIRE(c->n_ch_mult >= 1 && c->n_ch_mult <= 4 && c->num_res_blocks >= 1, "vqdec: bad ch_mult / num_res_blocks");
  EDB_REQUIRE(c->z_channels * 9 <= 64, "vqdec: z_channels*9 must fit one 64-wide K chunk");
  EDB_REQUIRE(cd->out_ch >= 1 && cd->out_ch <= 8 && c->embed_dim >= 1 && c->embed_dim <= 8, "vqdec: out_ch / embed_dim <= 8");
  EDB_REQUIRE(cd->disentangled_dim >= 0 && cd->disentangled_dim <= 64 && cd->n_embed >= 0, "vqdec: bad n_embed / disentangled_dim");
  const int L = c->n_ch_mult;
  const int R = c->resolution;
  EDB_REQUIRE(R >= 8 && R <= 128 && (R & (R - 1)) == 0, "vqdec: resolution must be a power of two in 8..128");
  const int r0 = R >> (L - 1);
  int block_in = c->ch * c->ch_mult[L - 1];
  EDB_REQUIRE(r0 >= 4 && r0 * r0 <= 256 && block_in <= 128 && block_in % 64 == 0,
              "vqdec: the mid attention kernel handles <= 256 tokens of 64 or 128 channels");
  u->vq_n_embed = cd->n_embed;
  u->vq_dd = cd->disentangled_dim;
  u->vq_img_ch = cd->out_ch;
  Builder b;
  b.u = u;
  auto conv_params = [&](const std::string& n, int co, int ci, int k, int& w, int& bi) {
    w = b.add_param(n + ".weight", {co, ci, k, k});
    bi = b.add_param(n + ".bias", {co});
  };
  auto norm_params = [&](const std::string& n, int ch, int& w, int& bi) {
    w = b.add_param(n + ".weight", {ch});
    bi = b.add_param(n + ".bias", {ch});
  };
  auto add_res = [&](const std::string& n, int ci, int co, int H) {
    VqLayer l;
    l.kind = 1;
    l.cin = ci;
    l.cout = co;
    l.H = H;
    norm_params(n + ".norm1", ci, l.n1w, l.n1b);
    conv_params(n + ".conv1", co, ci, 3, l.c1w, l.c1b);
    norm_params(n + ".norm2", co, l.n2w, l.n2b);
    conv_params(n + ".conv2", co, co, 3, l.c2w, l.c2b);
    if (ci != co) conv_params(n + ".nin_shortcut", co, ci, 1, l.sw, l.sb);
    const long long ci64 = ceil_to(ci, 64), co64 = ceil_to(co, 64);
    l.w1 = b.alloc(co, 9 * ci64);
    b.pack_fwd(l.w1, l.c1w, co, ci, 9, 0, 0);
    l.w2 = b.alloc(co, 9 * co64 + (l.sw >= 0 ? ci64 : 0));
    b.pack_fwd(l.w2, l.c2w, co, co, 9, 0, 0);
    if (l.sw >= 0) {
      b.pack_fwd(l.w2, l.sw, co, ci, 1, 0, static_cast<int>(9 * co64));
      l.bias_off = b.f32_off;
      b.f32_off += ceil_to(co, 64);
      b.entry(3, l.c2b, l.sb, l.bias_off, co, 1, 1, 0, 0, 0, 0, 0);
    }
    u->vq.push_back(l);
    return static_cast<int>(u->vq.size()) - 1;
  };
  conv_params("decoder.conv_in", block_in, c->z_channels, 3, u->vq_in_w, u->vq_in_b);
  std::vector<int> order;
  order.push_back(add_res("decoder.mid.block_1", block_in, block_in, r0));
  {
    VqLayer l;
    l.kind = 3;
    l.cin = l.cout = block_in;
    l.H = r0;
    const std::string n = "decoder.mid.attn_1";
    norm_params(n + ".norm", block_in, l.n1w, l.n1b);
    conv_params(n + ".q", block_in, block_in, 1, l.qw, l.qb);
    conv_params(n + ".k", block_in, block_in, 1, l.kw, l.kb);
    conv_params(n + ".v", block_in, block_in, 1, l.vw, l.vb);
    conv_params(n + ".proj_out", block_in, block_in, 1, l.pw, l.pb);
    const int C = block_in;
    l.w1 = b.alloc(3 * C, ceil_to(C, 64));
    const int src[3] = {l.qw, l.kw, l.vw}, bsrc[3] = {l.qb, l.kb, l.vb};
    l.bias_off = b.f32_off;
    b.f32_off += ceil_to(3 * C, 64);
    for (int i = 0; i < 3; ++i) {
      b.pack_fwd(l.w1, src[i], C, C, 1, i * C, 0);
      b.entry(2, bsrc[i], -1, l.bias_off, C, 1, 1, 0, 0, 0, 0, i * C);
    }
    l.w2 = b.alloc(C, ceil_to(C, 64));
    b.pack_fwd(l.w2, l.pw, C, C, 1, 0, 0);
    u->vq.push_back(l);
    order.push_back(static_cast<int>(u->vq.size()) - 1);
  }
  order.push_back(add_res("decoder.mid.block_2", block_in, block_in, r0));
  // channel / resolution of every level at the time it EXECUTES (deepest first)
  int lvl_in[4], lvl_res[4];
  {
    int bi = block_in, rr = r0;
    for (int lvl = L - 1; lvl >= 0; --lvl) {
      lvl_in[lvl] = bi;
      lvl_res[lvl] = rr;
      bi = c->ch * c->ch_mult[lvl];
      EDB_REQUIRE(bi % 32 == 0, "vqdec: channel counts must be multiples of 32");
      if (lvl != 0) rr *= 2;
    }
  }
  std::vector<std::vector<int>> lvl_layers(L);
  for (int lvl = 0; lvl < L; ++lvl) {          // registration order: up.0 first
    int bi = lvl_in[lvl];
    const int bo = c->ch * c->ch_mult[lvl];
    for (int i = 0; i <= c->num_res_blocks; ++i) {
      lvl_layers[lvl].push_back(add_res("decoder.up." + std::to_string(lvl) + ".block." + std::to_string(i), bi, bo, lvl_res[lvl]));
      bi = bo;
    }
    if (lvl != 0) {
      VqLayer l;
      l.kind = 4;
      l.cin = l.cout = bo;
      l.H = lvl_res[lvl];
      conv_params("decoder.up." + std::to_string(lvl) + ".upsample.conv", bo, bo, 3, l.c1w, l.c1b);
      l.w1 = b.alloc(bo, 9 * ceil_to(bo, 64));
      b.pack_fwd(l.w1, l.c1w, bo, bo, 9, 0, 0);
      u->vq.push_back(l);
      lvl_layers[lvl].push_back(static_cast<int>(u->vq.size()) - 1);
    }
  }
  for (int lvl = L - 1; lvl >= 0; --lvl)
    for (int i : lvl_layers[lvl]) order.push_back(i);
  u->vq_order = order;
  const int c_last = c->ch * c->ch_mult[0];
  norm_params("decoder.norm_out", c_last, u->vq_no_w, u->vq_no_b);
  conv_params("decoder.conv_out", cd->out_ch, c_last, 3, u->vq_co_w, u->vq_co_b);
  if (cd->n_embed > 0) u->vq_emb_w = b.add_param("quantize.embedding.weight", {cd->n_embed, c->embed_dim});
  u->vq_pq_w = b.add_param("post_quant_conv.weight", {c->z_channels, c->embed_dim + cd->disentangled_dim, 1, 1});
  u->vq_pq_b = b.add_param("post_quant_conv.bias", {c->z_channels});
  u->vq_out_res = R;
  u->vq_out_ch = c_last;
  u->stemf = b.alloc(block_in, 64);
  b.entry(4, u->vq_in_w, -1, u->stemf.off, block_in, c->z_channels, 9, 0, 0, 64, 0, 0);
  u->headf = b.alloc(cd->out_ch, 9 * ceil_to(c_last, 64));
  b.pack_fwd(u->headf, u->vq_co_w, cd->out_ch, c_last, 9, 0, 0);
  u->table_bytes = static_cast<size_t>(ceil_to(static_cast<long long>(u->table.size() * sizeof(PackEntry)), 1024));
  u->h16_bytes = static_cast<size_t>(ceil_to(b.bf_off * 2, 1024));
  u->f32_bytes = static_cast<size_t>(ceil_to(b.f32_off * 4, 1024));
  return EDB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Concept encoder trunk: Encoder4.encoder (openaimodel_enc.py:995-1014).  Parameters in registration order; operands:
// forward [Cout, taps * Cin] tap-major (the im2col column order), input-gradient [taps * Cin, Cout] (kind 6), the first
// layer in the reference's flat order over one 64-wide chunk.
// ---------------------------------------------------------------------------------------------------------------
int enc4_build(UNet* u, const encdiff_enc4_config* c) {
  u->model_kind = 3;
  u->e4.cfg = *c;
  memset(&u->cfg, 0, sizeof(u->cfg));
  EDB_REQUIRE(c->d == 128, "enc4: d must be 128 (the reference's View((-1, 128 * 4 * 4)) fixes it)");
  EDB_REQUIRE(c->resolution == 64, "enc4: resolution must be 64 (four stride-2 convolutions down to the 4 x 4 the Linear expects)");
  EDB_REQUIRE(c->num_channels * 16 <= 64, "enc4: num_channels * 16 must fit one 64-wide K chunk");
  EDB_REQUIRE(c->bn == 1, "enc4: only bn=True (the reference default) is implemented");
  EDB_REQUIRE(c->latent_unit >= 1 && c->latent_unit <= 64, "enc4: latent_unit must be in 1..64");
  UNet::E4& e = u->e4;
  Builder b;
  b.u = u;
  const int d = c->d, L = c->latent_unit;
  auto P = [&](const std::string& n, std::initializer_list<long long> shape) { return b.add_param("encoder." + n, shape); };
  const int conv_idx[4] = {0, 3, 6, 8}, bn_idx[4] = {1, 4, 7, 9};
  for (int i = 0; i < 4; ++i) {
    const int cin = i == 0 ? c->num_channels : d;
    e.cw[i] = P(std::to_string(conv_idx[i]) + ".weight", {d, cin, 4, 4});
    e.cb[i] = P(std::to_string(conv_idx[i]) + ".bias", {d});
    e.bw[i] = P(std::to_string(bn_idx[i]) + ".weight", {d});
    e.bb[i] = P(std::to_string(bn_idx[i]) + ".bias", {d});
  }
  auto resblock = [&](int r, int idx, int bn_slot) {
    const std::string n = std::to_string(idx) + ".convs.";
    e.r3w[r] = P(n + "1.weight", {d, d, 3, 3});
    e.r3b[r] = P(n + "1.bias", {d});
    e.bw[bn_slot] = P(n + "2.weight", {d});
    e.bb[bn_slot] = P(n + "2.bias", {d});
    e.r1w[r] = P(n + "4.weight", {d, d, 1, 1});
    e.r1b[r] = P(n + "4.bias", {d});
  };
  resblock(0, 11, 4);
  e.bw[5] = P("12.weight", {d});
  e.bb[5] = P("12.bias", {d});
  resblock(1, 14, 6);
  e.lw = P("16.weight", {L, 16LL * d});
  e.lb = P("16.bias", {L});
  // operands
  e.c0f = b.alloc(d, 64);
  b.entry(4, e.cw[0], -1, e.c0f.off, d, c->num_channels, 16, 0, 0, 64, 0, 0);
  for (int i = 1; i < 4; ++i) {
    e.cf[i] = b.alloc(d, 16LL * d);
    b.pack_fwd(e.cf[i], e.cw[i], d, d, 16, 0, 0);
    e.ct[i] = b.alloc(16LL * d, d);
    b.entry(6, e.cw[i], -1, e.ct[i].off, d, d, 16, d, 0, d, 0, 0);
  }
  for (int r = 0; r < 2; ++r) {
    e.r3f[r] = b.alloc(d, 9LL * d);
    b.pack_fwd(e.r3f[r], e.r3w[r], d, d, 9, 0, 0);
    e.r3t[r] = b.alloc(9LL * d, d);
    b.entry(6, e.r3w[r], -1, e.r3t[r].off, d, d, 9, d, 0, d, 0, 0);
    e.r1f[r] = b.alloc(d, d);
    b.pack_fwd(e.r1f[r], e.r1w[r], d, d, 1, 0, 0);
    e.r1t[r] = b.alloc(d, d);
    b.pack_t(e.r1t[r], e.r1w[r], d, d, 1, 0, 0);
  }
  e.lf = b.alloc(L, 16LL * d);
  b.pack_fwd(e.lf, e.lw, L, 16 * d, 1, 0, 0);
  e.lt = b.alloc(16LL * d, 64);
  b.pack_t(e.lt, e.lw, L, 16 * d, 1, 0, 0);
  u->table_bytes = static_cast<size_t>(ceil_to(static_cast<long long>(u->table.size() * sizeof(PackEntry)), 1024));
  u->h16_bytes = static_cast<size_t>(ceil_to(b.bf_off * 2, 1024));
  u->f32_bytes = static_cast<size_t>(ceil_to(b.f32_off * 4, 1024));
  return EDB_OK;
}

}  // namespace edb
