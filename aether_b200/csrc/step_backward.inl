// Backward of the step (included by step.cu).  Reverse of forward(): every Linear contributes a
// data gradient (GEMM_NN, fused silu' / mask epilogue), a weight gradient (GEMM_TN, deterministic
// split-K) and a bias gradient (column sums); the receiver mean becomes a broadcast over the CSR
// segment; the two gathers of the node-factorised first Linear become segment sums over the
// receiver- and sender-grouped views -- no atomics anywhere.
namespace aether {
namespace {

struct Bwd {
  const AetherDims& d; const float* const* P; float* const* G; const Scratch& s; cudaStream_t st; NodeTc nc;

  // dW[N_out, K_in] = gy^T act(a)   ;   db = colsum(gy)
  int wgrad(const float* gy, int64_t ldgy, int64_t rows, int n_out, const float* a, int64_t lda, int k_in, int tf,
            const float* mask, float* dW, int64_t lddw, float* db) const {
    const int rc = wgrad_tc(nc, gy, ldgy, rows, n_out, a, lda, k_in, tf, mask, dW, lddw, st);
    if (rc != AETHER_E_UNSUPPORTED) {
      if (rc != 0) return rc;
      if (db) AE_CHECK(colsum_f32(gy, ldgy, rows, n_out, db, s.partial, s.partial_floats, st));
      return 0;
    }
    GemmArgs g;
    g.A = gy; g.lda = ldgy; g.B = a; g.ldb = lda; g.C = dW; g.ldc = lddw; g.M = n_out; g.N = k_in; g.K = rows;
    g.tf = tf; g.tf_mask = mask; g.partial = s.partial; g.partial_floats = s.partial_floats;
    AE_CHECK(gemm_f32(GEMM_TN, g, st));
    if (db) AE_CHECK(colsum_f32(gy, ldgy, rows, n_out, db, s.partial, s.partial_floats, st));
    return 0;
  }
  // gx[rows, K_in] = epi(gy W)    W is [N_out, K_in] with leading dimension ldw
  int dgrad(const float* gy, int64_t ldgy, int64_t rows, int n_out, const float* W, int64_t ldw, int k_in, float* gx,
            int64_t ldgx, const float* dsilu_src, const float* mask, const float* res, int accumulate) const {
    const int rc = dgrad_tc(nc, gy, ldgy, rows, n_out, W, ldw, k_in, gx, ldgx, dsilu_src, mask, res, accumulate, st);
    if (rc != AETHER_E_UNSUPPORTED) return rc;
    GemmArgs g;
    g.A = gy; g.lda = ldgy; g.B = W; g.ldb = ldw; g.C = gx; g.ldc = ldgx; g.M = rows; g.N = k_in; g.K = n_out;
    g.dsilu = dsilu_src; g.ldd = ldgx; g.emask = mask; g.ldm = ldgx; g.res = res; g.ldr = ldgx;
    g.accumulate = accumulate;
    return gemm_f32(GEMM_NN, g, st);
  }
};

// Edge part of the backward pass on the bf16 tensor-core path, interleaved with the node-level callbacks.
// Per layer l (top down), with g2_l = dL/dpre2_l already in s.bg2:
//   B1   g1 = (g2 W2) * silu'(pre1_l)                                   tc_gemm, silu' epilogue
//   dW2 = g2^T silu(pre1_l), db2 ; dW1e = g1^T silu(pre2_{l-1}) (layer 1: g1^T feat), db1       tc_wgrad
//   l > 1: segment sums of g1 by sender / receiver -> node terms -> grad of x_{l-1} -> node update of
//          layer l-1 -> gy_{l-1}; then ONE GEMM produces the next g2 directly:
//   B2   g2_{l-1} = (g1 We + gy_{l-1}[recv] / deg) * silu'(pre2_{l-1})   gather + silu' epilogue
// so dL/de_{l-1} is never written to HBM and the separate elementwise pass exists only for the top layer.
template <class FA, class FP, class FR>
int backward_edges_tc(const AetherDims& d, const float* const* P, const AetherGraph& Gr, float* const* G,
                      const Workspace& w, const Scratch& s, cudaStream_t st, FA& node_update_bwd, FP& node_terms_bwd,
                      FR& node_res_bwd) {
  const int D = d.num_dims, H = d.hidden, Fe = num_edge_feat(D);
  const int64_t Nn = d.n_nodes, E = d.n_edges;
  void* g2 = s.bg2;
  void* g1 = s.bg1;
  void* g1_next = s.bg1b;       // the chained kernel writes the next layer's g1 while it reads this layer's
  const bool chain = use_chain() && tc_chain_bwd_supported(H);
  bool have_g1 = false;         // g1 of this layer was already produced by the previous iteration's chained launch
  AE_CHECK(node_update_bwd(4));
  if (chain) {
    // top layer: g2 = gyd[recv] * silu'(pre2_4) and g1 = (g2 W2) * silu'(pre1_4) in one launch of the chained kernel
    // without its first GEMM (no standalone elementwise pass, no re-read of g2)
    ProfScope ps(PROF_EDGE_BWD_DATA, st);
    AE_CHECK(launch_node_mean_scale_bf16(Nn, H, Gr.rowptr, s.gy, s.gyd16, st));
    TcChainArgs c;
    c.M = E; c.wpack2 = w.wtB[4];
    c.gat16 = (const __nv_bfloat16*)s.gyd16; c.ldg = H; c.idx_b = Gr.recv;
    c.out1 = (__nv_bfloat16*)g2; c.ldo1 = H; c.out2 = (__nv_bfloat16*)g1; c.ldo2 = H;
    AE_CHECK(tc_chain_bwd(c, nullptr, 0, w.bpre2[4], H, w.bpre1[4], H, H, st));
    have_g1 = true;
  } else {
    ProfScope ps(PROF_EDGE_BWD_DATA, st);
    AE_CHECK(launch_edge_dpre2_bf16(E, H, Gr.recv, Gr.rowptr, nullptr, s.gy, w.bpre2[4], g2, st));
  }
  for (int l = 4; l >= 1; --l) {
    const LayerSlots sl = layer_slots(l);
    if (!have_g1) {
      ProfScope ps(PROF_EDGE_BWD_DATA, st);
      TcGemmArgs a;
      a.M = E; a.wpack = w.wtB[l]; a.dsilu = (const __nv_bfloat16*)w.bpre1[l]; a.ldd = H;
      a.out_bf16 = (__nv_bfloat16*)g1; a.ldo = H;
      AE_CHECK(tc_gemm(a, g2, H, H, H, TC_TF_NONE, st));
    }
    {
      ProfScope ps(PROF_EDGE_WGRAD, st);
      AE_CHECK(tc_wgrad(g2, H, H, w.bpre1[l], H, H, TC_TF_SILU, E, G[sl.m2w], H, H, 0, G[sl.m2b], s.partial,
                        s.partial_floats, st));
      if (l > 1)
        AE_CHECK(tc_wgrad(g1, H, H, w.bpre2[l - 1], H, H, TC_TF_SILU, E, G[sl.m0w] + 2 * H, 3 * H, H, 0, G[sl.m0b],
                          s.partial, s.partial_floats, st));
      else
        AE_CHECK(tc_wgrad(g1, H, H, w.feat16, 64, 64, TC_TF_NONE, E, G[sl.m0w], Fe, Fe, 1, G[sl.m0b], s.partial,
                          s.partial_floats, st));
    }
    if (l > 1) {
      {
        ProfScope ps(PROF_SEGMENT, st);
        __nv_bfloat16* gp16 = (__nv_bfloat16*)s.gpspr16;     // bf16 copies: operands of the node terms' tensor-core GEMMs
        AE_CHECK(launch_segsum_bf16(Nn, H, Gr.colptr, Gr.sperm, g1, H, s.gpspr, 2 * H, st, gp16, 2 * H));
        AE_CHECK(launch_segsum_bf16(Nn, H, Gr.rowptr, nullptr, g1, H, s.gpspr + H, 2 * H, st, gp16 + H, 2 * H));
      }
      AE_CHECK(node_terms_bwd(l));
      AE_CHECK(node_update_bwd(l - 1));
      ProfScope ps(PROF_EDGE_BWD_DATA, st);
      AE_CHECK(launch_node_mean_scale_bf16(Nn, H, Gr.rowptr, s.gy, s.gyd16, st));
      if (chain) {
        // B2 of this layer and B1 of the next one down in ONE launch (tc_chain_bwd.cu): g2_{l-1} is stored for the weight
        // gradient and handed to the second GEMM through tensor memory
        TcChainArgs c;
        c.M = E; c.wpack1 = w.wtA[l]; c.wpack2 = w.wtB[l - 1];
        c.gat16 = (const __nv_bfloat16*)s.gyd16; c.ldg = H; c.idx_b = Gr.recv;
        c.out1 = (__nv_bfloat16*)g2; c.ldo1 = H; c.out2 = (__nv_bfloat16*)g1_next; c.ldo2 = H;
        AE_CHECK(tc_chain_bwd(c, g1, H, w.bpre2[l - 1], H, w.bpre1[l - 1], H, H, st));
        void* t = g1; g1 = g1_next; g1_next = t;
        have_g1 = true;
        continue;
      }
      TcGemmArgs a;
      a.M = E; a.wpack = w.wtA[l];
      a.gat16 = (const __nv_bfloat16*)s.gyd16; a.ldg = H; a.idx_a = Gr.recv;
      a.dsilu = (const __nv_bfloat16*)w.bpre2[l - 1]; a.ldd = H;
      a.out_bf16 = (__nv_bfloat16*)g2; a.ldo = H;
      AE_CHECK(tc_gemm(a, g1, H, H, H, TC_TF_NONE, st));
    } else {
      // gfeat[E, 32] = g1 W1 (fp32, consumed by the geometry backward)
      TcGemmArgs a;
      a.M = E; a.wpack = w.wtA[1]; a.out_f32 = s.gfeat; a.ldo32 = 32; a.n_valid = 32;
      AE_CHECK(tc_gemm(a, g1, H, H, 32, TC_TF_NONE, st));
      AE_CHECK(node_res_bwd());
    }
  }
  return 0;
}

int backward(const AetherDims& d, const float* const* P, const AetherGraph& Gr, const float* x, const float* vel,
             const float* q, const float* dmask, const float* gout, float* const* G, float* gx_out, float* gvel_out,
             const Workspace& w, const Scratch& s, uint32_t flags, cudaStream_t st) {
  const int D = d.num_dims, H = d.hidden, Fe = num_edge_feat(D), ldf = edge_feat_ld(D);
  const int64_t Nn = d.n_nodes, E = d.n_edges;
  NodeTc nc;
  nc.on = use_tc(d, flags); nc.a16 = s.na16; nc.g16 = s.ng16; nc.wp = s.nwp; nc.partial = s.partial;
  nc.partial_floats = s.partial_floats;
  const Bwd b{d, P, G, s, st, nc};
  const float* m1 = dmask;
  const float* m2 = dmask ? dmask + (size_t)Nn * H : nullptr;

  // out = x + R pred
  std::unique_ptr<ProfScope> pn(new ProfScope(PROF_NODE, st));   // node-level head of the backward (closed before the edge part)
  AE_CHECK(launch_globalize_bwd(D, Nn, w.R, w.pred, gout, s.gpred, s.gRglob, st));
  // out MLP
  AE_CHECK(b.wgrad(s.gpred, D, Nn, D, w.o2, H, H, m2 ? TF_SILU_MASK : TF_SILU, m2, G[P_OUT6_W], H, G[P_OUT6_B]));
  AE_CHECK(b.dgrad(s.gpred, D, Nn, D, P[P_OUT6_W], H, H, s.go2, H, w.o2, m2, nullptr, 0));
  AE_CHECK(b.wgrad(s.go2, H, Nn, H, w.o1, H, H, m1 ? TF_SILU_MASK : TF_SILU, m1, G[P_OUT3_W], H, G[P_OUT3_B]));
  AE_CHECK(b.dgrad(s.go2, H, Nn, H, P[P_OUT3_W], H, H, s.go1, H, w.o1, m1, nullptr, 0));
  AE_CHECK(b.wgrad(s.go1, H, Nn, H, w.xl[4], H, H, TF_NONE, nullptr, G[P_OUT0_W], H, G[P_OUT0_B]));
  float* gx = s.gx;      // gradient w.r.t. the current layer's output node state
  float* gxn = s.gx2;    // ... and the next one down
  AE_CHECK(b.dgrad(s.go1, H, Nn, H, P[P_OUT0_W], H, H, gx, H, nullptr, nullptr, nullptr, 0));
  const float* ge = nullptr;   // gradient w.r.t. e_l from layer l+1 (none for the last layer)
  const bool tc = use_tc(d, flags);
  const int ldgf = tc ? 32 : ldf;   // leading dimension of gfeat
  pn.reset();

  // node part A of layer l: x_l = y + W4 silu(pre3) + b4 ; pre3 = W3 y + b3   (gx -> s.gy)
  void* gpre316 = s.gpre316;
  auto node_update_bwd = [&](int l) -> int {
    ProfScope pn(PROF_NODE, st);
    const LayerSlots sl = layer_slots(l);
    if (tc && use_chain()) {
      // bf16 operands throughout (forward kept y_l and pre3_l as bf16, tc_chain.cu): no conversion passes except gx
      const __nv_bfloat16* p316 = (const __nv_bfloat16*)w.pre316[l];
      const __nv_bfloat16* y16 = (const __nv_bfloat16*)w.y16[l];
      __nv_bfloat16* g316 = (__nv_bfloat16*)gpre316;
      AE_CHECK(tc_convert_bf16(gx, H, Nn, H, TF_NONE, nullptr, s.na16, H, st));
      for (int h = 0; h < 2; ++h) {
        // dW4[:, half] = gx^T silu(pre3[:, half]) ; db4 = colsum(gx)
        AE_CHECK(tc_wgrad(s.na16, H, H, p316 + h * H, 2 * H, H, TC_TF_SILU, Nn, G[sl.u2w] + h * H, 2 * H, H, 0,
                          h == 0 ? G[sl.u2b] : nullptr, s.partial, s.partial_floats, st));
        // gpre3[:, half] = (gx W4[:, half]) * silu'(pre3[:, half])
        TcGemmArgs a;
        a.M = Nn; a.wpack = w.nt4[l][h]; a.dsilu = p316 + h * H; a.ldd = 2 * H; a.out_bf16 = g316 + h * H; a.ldo = 2 * H;
        AE_CHECK(tc_gemm(a, s.na16, H, H, H, TC_TF_NONE, st));
      }
      for (int h = 0; h < 2; ++h) {
        // dW3[half, :] = gpre3[:, half]^T y ; db3[half] = colsum
        AE_CHECK(tc_wgrad(g316 + h * H, 2 * H, H, y16, H, H, TC_TF_NONE, Nn, G[sl.u0w] + (size_t)h * H * H, H, H, 0,
                          G[sl.u0b] + h * H, s.partial, s.partial_floats, st));
        // gy = gx + gpre3 W3   (two K halves, the second accumulates)
        TcGemmArgs a;
        a.M = Nn; a.wpack = w.nt3[l][h]; a.res = h == 0 ? gx : s.gy; a.ldr = H; a.out_f32 = s.gy; a.ldo32 = H; a.n_valid = H;
        AE_CHECK(tc_gemm(a, g316 + h * H, 2 * H, H, H, TC_TF_NONE, st));
      }
      return 0;
    }
    AE_CHECK(b.wgrad(gx, H, Nn, H, w.pre3[l], 2 * H, 2 * H, TF_SILU, nullptr, G[sl.u2w], 2 * H, G[sl.u2b]));
    AE_CHECK(b.dgrad(gx, H, Nn, H, P[sl.u2w], 2 * H, 2 * H, s.gpre3, 2 * H, w.pre3[l], nullptr, nullptr, 0));
    AE_CHECK(b.wgrad(s.gpre3, 2 * H, Nn, 2 * H, w.y[l], H, H, TF_NONE, nullptr, G[sl.u0w], H, G[sl.u0b]));
    AE_CHECK(b.dgrad(s.gpre3, 2 * H, Nn, 2 * H, P[sl.u0w], H, H, s.gy, H, nullptr, nullptr, gx, 0));
    return 0;
  };
  // node terms of layer l > 1: P = x_{l-1} [Ws; Wr]^T from the segment sums s.gpspr ; gx <- grad of x_{l-1}
  auto node_terms_bwd = [&](int l) -> int {
    ProfScope pn(PROF_NODE, st);
    const LayerSlots sl = layer_slots(l);
    const float* xin = w.xl[l - 1];
    if (tc && use_chain()) {
      // bf16 operands from the producers (segment sums wrote gpspr16, the forward's node update wrote x16): no
      // conversion or packing launches
      const __nv_bfloat16* gp16 = (const __nv_bfloat16*)s.gpspr16;
      const __nv_bfloat16* x16 = (const __nv_bfloat16*)w.x16[l - 1];
      AE_CHECK(tc_wgrad(gp16, 2 * H, H, x16, H, H, TC_TF_NONE, Nn, G[sl.m0w], 3 * H, H, 0, nullptr, s.partial, s.partial_floats, st));
      AE_CHECK(tc_wgrad(gp16 + H, 2 * H, H, x16, H, H, TC_TF_NONE, Nn, G[sl.m0w] + H, 3 * H, H, 0, nullptr, s.partial,
                        s.partial_floats, st));
      TcGemmArgs a;
      a.M = Nn; a.wpack = w.nts[l]; a.res = s.gy; a.ldr = H; a.out_f32 = gxn; a.ldo32 = H; a.n_valid = H;
      AE_CHECK(tc_gemm(a, gp16, 2 * H, H, H, TC_TF_NONE, st));
      TcGemmArgs c;
      c.M = Nn; c.wpack = w.ntr[l]; c.accumulate = 1; c.out_f32 = gxn; c.ldo32 = H; c.n_valid = H;
      AE_CHECK(tc_gemm(c, gp16 + H, 2 * H, H, H, TC_TF_NONE, st));
      float* t = gx; gx = gxn; gxn = t;
      return 0;
    }
    AE_CHECK(b.wgrad(s.gpspr, 2 * H, Nn, H, xin, H, H, TF_NONE, nullptr, G[sl.m0w], 3 * H, nullptr));
    AE_CHECK(b.wgrad(s.gpspr + H, 2 * H, Nn, H, xin, H, H, TF_NONE, nullptr, G[sl.m0w] + H, 3 * H, nullptr));
    AE_CHECK(b.dgrad(s.gpspr, 2 * H, Nn, H, P[sl.m0w], 3 * H, H, gxn, H, nullptr, nullptr, s.gy, 0));
    AE_CHECK(b.dgrad(s.gpspr + H, 2 * H, Nn, H, P[sl.m0w] + H, 3 * H, H, gxn, H, nullptr, nullptr, nullptr, 1));
    float* t = gx; gx = gxn; gxn = t;
    return 0;
  };
  // layer 1: y1 = Wres x0 + bres + agg
  auto node_res_bwd = [&]() -> int {
    ProfScope pn(PROF_NODE, st);
    AE_CHECK(b.wgrad(s.gy, H, Nn, H, w.x0, 3 * D, 3 * D, TF_NONE, nullptr, G[P_L1_RES_W], 3 * D, G[P_L1_RES_B]));
    AE_CHECK(b.dgrad(s.gy, H, Nn, H, P[P_L1_RES_W], 3 * D, 3 * D, s.gx0, 3 * D, nullptr, nullptr, nullptr, 0));
    return 0;
  };

  if (tc) {
    AE_CHECK(backward_edges_tc(d, P, Gr, G, w, s, st, node_update_bwd, node_terms_bwd, node_res_bwd));
  } else {
    for (int l = 4; l >= 1; --l) {
      const LayerSlots sl = layer_slots(l);
      AE_CHECK(node_update_bwd(l));
      // y = base + mean_recv(silu(pre2)) : broadcast gy / deg over the segment, add grad from layer l+1
      float* g2 = s.gA;
      float* g1 = s.gB;
      {
        ProfScope ps(PROF_EDGE_BWD_DATA, st);
        AE_CHECK(launch_edge_dpre2(E, H, Gr.recv, Gr.rowptr, ge, s.gy, w.pre2[l], g2, st));
        // pre2 = W2 silu(pre1) + b2
        AE_CHECK(b.dgrad(g2, H, E, H, P[sl.m2w], H, H, g1, H, w.pre1[l], nullptr, nullptr, 0));
      }
      {
        ProfScope ps(PROF_EDGE_WGRAD, st);
        AE_CHECK(b.wgrad(g2, H, E, H, w.pre1[l], H, H, TF_SILU, nullptr, G[sl.m2w], H, G[sl.m2b]));
        AE_CHECK(colsum_f32(g1, H, E, H, G[sl.m0b], s.partial, s.partial_floats, st));
      }
      if (l > 1) {
        // pre1 = Ps[send] + Pr[recv] + We silu(pre2_{l-1}) + b1
        {
          ProfScope ps(PROF_EDGE_WGRAD, st);
          AE_CHECK(b.wgrad(g1, H, E, H, w.pre2[l - 1], H, H, TF_SILU, nullptr, G[sl.m0w] + 2 * H, 3 * H, nullptr));
        }
        {
          ProfScope ps(PROF_EDGE_BWD_DATA, st);
          AE_CHECK(b.dgrad(g1, H, E, H, P[sl.m0w] + 2 * H, 3 * H, H, s.gA, H, nullptr, nullptr, nullptr, 0));
        }
        ge = s.gA;
        {
          ProfScope ps(PROF_SEGMENT, st);
          AE_CHECK(launch_segsum(Nn, H, Gr.colptr, Gr.sperm, g1, H, s.gpspr, 2 * H, 0, st));
          AE_CHECK(launch_segsum(Nn, H, Gr.rowptr, nullptr, g1, H, s.gpspr + H, 2 * H, 0, st));
        }
      } else {
        // pre1 = W1 feat + b1
        AE_CHECK(b.wgrad(g1, H, E, H, w.feat, ldf, Fe, TF_NONE, nullptr, G[sl.m0w], Fe, nullptr));
        GemmArgs g;
        g.A = g1; g.lda = H; g.B = P[sl.m0w]; g.ldb = Fe; g.C = s.gfeat; g.ldc = ldf; g.M = E; g.N = Fe; g.K = H;
        AE_CHECK(gemm_f32(GEMM_NN, g, st));
      }
      if (l > 1) AE_CHECK(node_terms_bwd(l));
      else AE_CHECK(node_res_bwd());
    }
  }

  // localizer: per-edge partials, then segment sums by receiver (contiguous) and by sender (sperm)
  ProfScope pn2(PROF_NODE, st);                      // geometry + field backward: counted with the node-level work
  const int GS = geom_gs(D), GR = geom_gr(D);
  AE_CHECK(launch_edge_features_bwd(D, E, Gr.send, Gr.recv, x, vel, w.f, w.R, s.gfeat, ldgf, s.gsend, s.grecv, st));
  AE_CHECK(launch_segsum(Nn, GS, Gr.colptr, Gr.sperm, s.gsend, GS, s.ssend, GS, 0, st));
  AE_CHECK(launch_segsum(Nn, GR, Gr.rowptr, nullptr, s.grecv, GR, s.srecv, GR, 0, st));
  float* gxin = gx_out ? gx_out : s.gxin;
  float* gvin = gvel_out ? gvel_out : s.gvin;
  AE_CHECK(launch_node_frame_bwd(D, Nn, vel, w.f, w.R, s.gx0, s.ssend, s.srecv, s.gRglob, gout, s.gf, gxin, gvin, st));
  if (flags & AETHER_F_EXTERNAL_FIELD) {
    // the field was an input: hand its gradient to the caller (dynfield.cu continues from there)
    return (int)cudaMemcpyAsync(G[P_F0_W], s.gf, (size_t)Nn * D * sizeof(float), cudaMemcpyDeviceToDevice, st);
  }
  // field network (aether.py:113-119,127-134): input gradient and all six parameter gradients in two launches (field.cu)
  AE_CHECK(launch_field_mlp_bwd(D, Nn, s.gf, w.fin, w.fz1, w.fz2, P[P_F0_W], P[P_F2_W], P[P_F4_W], s.gfin, G[P_F0_W], G[P_F0_B],
                                G[P_F2_W], G[P_F2_B], G[P_F4_W], G[P_F4_B], s.partial, s.partial_floats, st));
  AE_CHECK(launch_embedding_bwd(D, Nn, q, s.gfin, G[P_EMB], s.partial, s.partial_floats, st));
  if (gx_out || gvel_out) AE_CHECK(launch_add_field_input_grads(D, Nn, s.gfin, gx_out, gvel_out, st));
  return 0;
}

}  // namespace
}  // namespace aether
