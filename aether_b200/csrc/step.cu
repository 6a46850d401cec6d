// Host-side orchestration of the Aether state-to-state step and its C ABI.
//   forward : nn/state2state/aether.py:169-186
//   backward: autograd of the same (experiments/lorentz/main.py:291)
// Every launch goes to the caller's stream; nothing synchronises.
#include <stdlib.h>
#include <string.h>

#include <memory>

#include "common.cuh"
#include "param_slots.cuh"
#include "tc_gemm.cuh"

namespace aether {

namespace {

struct Carver {
  char* base; size_t off = 0;
  explicit Carver(void* p) : base((char*)p) {}
  float* f(size_t n) { float* r = (float*)(base + off); off += align_up(n * sizeof(float), 256); return r; }
  void* b(size_t bytes) { void* r = base + off; off += align_up(bytes, 1024); return r; }
};

// bf16 tensor-core path for the edge MLPs: requested by the flag and available for this hidden size
inline bool use_tc(const AetherDims& d, uint32_t flags) {
  return (flags & AETHER_F_BF16_EDGE_MLP) && (d.hidden == 64 || d.hidden == 128 || d.hidden == 256);
}

// Everything the forward pass produces.  With AETHER_F_SAVE_FOR_BACKWARD each layer keeps its own
// edge pre-activations; otherwise pre1 is shared and pre2 ping-pongs between two buffers.
struct Workspace {
  float *fin, *fz1, *fz2, *f, *R, *x0, *feat;
  float *pre1[5], *pre2[5];
  float *pspr, *y[5], *pre3[5], *xl[5];
  float *o1, *o2, *pred;
  // bf16 path: features [E,64] (hi|lo), pre-activations [E,H] bf16, packed weight images per layer
  void* feat16; void *bpre1[5], *bpre2[5]; void *wpA[5], *wpB[5];
  void *na16, *nwp;     // node-level tensor-core scratch: bf16 operand copy [Nn,256], weight image
  void* pspr16;         // [Nn, 2H] bf16 node terms P_s | P_r of the factorised first Linear
  // node update on the chained kernel: bf16 operand copies y_l, x_l [Nn,H], pre3_l [Nn,2H] and the packed images of
  // W3 / W4 (two H x H halves each) and of the next layer's W_s, W_r
  void *y16[5], *x16[5], *pre316[5], *nw3[5][2], *nw4[5][2], *nws[5], *nwr[5];
  // transposed images for the backward (packed by the forward's single batched launch when it saves for backward)
  void *wtA[5], *wtB[5], *nt3[5][2], *nt4[5][2], *nts[5], *ntr[5];   // nts / ntr: W_s^T, W_r^T of layer l's first Linear
  size_t bytes;
};

Workspace carve(const AetherDims& d, uint32_t flags, void* base) {
  Workspace w;
  Carver c(base);
  const size_t Nn = (size_t)d.n_nodes, E = (size_t)d.n_edges, D = d.num_dims, H = d.hidden;
  const bool save = flags & AETHER_F_SAVE_FOR_BACKWARD;
  w.fin = c.f(Nn * (2 * D + 16));
  w.fz1 = c.f(Nn * 32);
  w.fz2 = c.f(Nn * 32);
  w.f = c.f(Nn * D);
  w.R = c.f(Nn * D * D);
  w.x0 = c.f(Nn * 3 * D);
  w.pre1[0] = w.pre2[0] = nullptr;
  w.feat = nullptr; w.feat16 = nullptr; w.na16 = w.nwp = w.pspr16 = nullptr;
  for (int l = 0; l <= 4; ++l) {
    w.bpre1[l] = w.bpre2[l] = w.wpA[l] = w.wpB[l] = nullptr; w.pre1[l] = w.pre2[l] = nullptr;
    w.y16[l] = w.x16[l] = w.pre316[l] = w.nws[l] = w.nwr[l] = nullptr;
    w.nw3[l][0] = w.nw3[l][1] = w.nw4[l][0] = w.nw4[l][1] = nullptr;
    w.wtA[l] = w.wtB[l] = w.nt3[l][0] = w.nt3[l][1] = w.nt4[l][0] = w.nt4[l][1] = w.nts[l] = w.ntr[l] = nullptr;
  }
  if (use_tc(d, flags)) {
    w.na16 = c.b(Nn * 256 * 2);
    w.nwp = c.b(tc_weight_pack_bytes(256, 256));
    w.pspr16 = c.b(Nn * 2 * H * 2);
    w.feat16 = c.b(E * 64 * 2);
    if (save) {
      for (int l = 1; l <= 4; ++l) { w.bpre1[l] = c.b(E * H * 2); w.bpre2[l] = c.b(E * H * 2); }
    } else {
      void* p1 = c.b(E * H * 2);
      void* a = c.b(E * H * 2);
      void* b = c.b(E * H * 2);
      for (int l = 1; l <= 4; ++l) { w.bpre1[l] = p1; w.bpre2[l] = (l & 1) ? a : b; }
    }
    for (int l = 1; l <= 4; ++l) { w.wpA[l] = c.b(tc_weight_pack_bytes(l == 1 ? 64 : H, H)); w.wpB[l] = c.b(tc_weight_pack_bytes(H, H)); }
    for (int l = 1; l <= 4; ++l) {
      w.y16[l] = c.b(Nn * H * 2); w.x16[l] = c.b(Nn * H * 2); w.pre316[l] = c.b(Nn * 2 * H * 2);
      for (int h = 0; h < 2; ++h) { w.nw3[l][h] = c.b(tc_weight_pack_bytes(H, H)); w.nw4[l][h] = c.b(tc_weight_pack_bytes(H, H)); }
      w.nws[l] = c.b(tc_weight_pack_bytes(H, H)); w.nwr[l] = c.b(tc_weight_pack_bytes(H, H));
    }
    if (save) {
      for (int l = 1; l <= 4; ++l) {
        w.wtA[l] = c.b(tc_weight_pack_bytes(H, H)); w.wtB[l] = c.b(tc_weight_pack_bytes(H, H));
        for (int h = 0; h < 2; ++h) { w.nt3[l][h] = c.b(tc_weight_pack_bytes(H, H)); w.nt4[l][h] = c.b(tc_weight_pack_bytes(H, H)); }
        if (l > 1) { w.nts[l] = c.b(tc_weight_pack_bytes(H, H)); w.ntr[l] = c.b(tc_weight_pack_bytes(H, H)); }
      }
    }
  } else {
  w.feat = c.f(E * edge_feat_ld(d.num_dims));
  if (save) {
    for (int l = 1; l <= 4; ++l) { w.pre1[l] = c.f(E * H); w.pre2[l] = c.f(E * H); }
  } else {
    float* p1 = c.f(E * H);
    float* a = c.f(E * H);
    float* b = c.f(E * H);
    for (int l = 1; l <= 4; ++l) { w.pre1[l] = p1; w.pre2[l] = (l & 1) ? a : b; }
  }
  }
  w.pspr = c.f(Nn * 2 * H);
  w.y[0] = w.pre3[0] = nullptr;
  w.xl[0] = w.x0;
  for (int l = 1; l <= 4; ++l) { w.y[l] = c.f(Nn * H); w.pre3[l] = c.f(Nn * 2 * H); w.xl[l] = c.f(Nn * H); }
  w.o1 = c.f(Nn * H);
  w.o2 = c.f(Nn * H);
  w.pred = c.f(Nn * D);
  w.bytes = c.off;
  return w;
}

struct Scratch {
  float *gA, *gB, *gfeat, *gsend, *grecv;
  float *gx, *gx2, *gy, *gpre3, *gpspr, *gx0, *go1, *go2, *gpred, *gRglob, *ssend, *srecv;
  float *gf, *gfz1, *gfz2, *gfin, *gxin, *gvin;
  float* partial; size_t partial_floats;
  // bf16 path: edge gradients [E,H] bf16 (ge, g2, g1), transposed weight images
  void *bg2, *bg1, *bg1b, *gyd16; void *wtA, *wtB;   // gyd16: [Nn,H] bf16 = gy / in-degree
  void *na16, *ng16, *nwp;
  void* gpre316;        // [Nn, 2H] bf16 gradient of the node update's hidden layer
  void* gpspr16;        // [Nn, 2H] bf16 copy of the node terms' gradients (segment sums of g1)
  size_t bytes;
};

size_t partial_need(const AetherDims& d, uint32_t flags) {
  const int64_t Nn = d.n_nodes, E = d.n_edges;
  const int D = d.num_dims, H = d.hidden, Fe = num_edge_feat(D);
  size_t m = 0;
  auto up = [&](size_t v) { if (v > m) m = v; };
  if (use_tc(d, flags)) {
    const int c2 = 2 * H > 256 ? 256 : 2 * H;
    up(tc_wgrad_partial_floats(c2, c2));
    up(tc_wgrad_partial_floats(H, 64));
    up(colsum_bf16_partial_floats(E, H));
  } else {
  up(gemm_tn_partial_floats(H, H, E));
  up(gemm_tn_partial_floats(H, Fe, E));
  up(colsum_partial_floats(E, H));
  }
  up(gemm_tn_partial_floats(2 * H, H, Nn));
  up(gemm_tn_partial_floats(H, 2 * H, Nn));
  up(gemm_tn_partial_floats(H, H, Nn));
  up(gemm_tn_partial_floats(H, 3 * D, Nn));
  up(gemm_tn_partial_floats(D, H, Nn));
  up(gemm_tn_partial_floats(32, 2 * D + 16, Nn));
  up(gemm_tn_partial_floats(32, 32, Nn));
  up(gemm_tn_partial_floats(D, 32, Nn));
  up(colsum_partial_floats(Nn, 2 * H));
  up(256 * 48);   // embedding gradient: per-block class sums
  up(field_mlp_bwd_partial_floats(D, Nn));
  return m + 64;
}

Scratch carve_scratch(const AetherDims& d, uint32_t flags, void* base) {
  Scratch s;
  Carver c(base);
  const size_t Nn = (size_t)d.n_nodes, E = (size_t)d.n_edges, D = d.num_dims, H = d.hidden;
  s.gA = s.gB = nullptr;
  s.bg2 = s.bg1 = s.bg1b = s.gyd16 = s.wtA = s.wtB = nullptr;
  s.na16 = s.ng16 = s.nwp = nullptr; s.gpre316 = s.gpspr16 = nullptr;
  if (use_tc(d, flags)) {
    s.na16 = c.b(Nn * 256 * 2); s.ng16 = c.b(Nn * 256 * 2); s.nwp = c.b(tc_weight_pack_bytes(256, 256));
    s.gpre316 = c.b(Nn * 2 * H * 2); s.gpspr16 = c.b(Nn * 2 * H * 2);
    s.bg2 = c.b(E * H * 2); s.bg1 = c.b(E * H * 2); s.bg1b = c.b(E * H * 2); s.gyd16 = c.b(Nn * H * 2);
    s.wtA = c.b(tc_weight_pack_bytes(H, H)); s.wtB = c.b(tc_weight_pack_bytes(H, H));
    s.gfeat = c.f(E * 32);
  } else {
    s.gA = c.f(E * H);
    s.gB = c.f(E * H);
    s.gfeat = c.f(E * edge_feat_ld(d.num_dims));
  }
  s.gsend = c.f(E * geom_gs(d.num_dims));
  s.grecv = c.f(E * geom_gr(d.num_dims));
  s.gx = c.f(Nn * H); s.gx2 = c.f(Nn * H); s.gy = c.f(Nn * H);
  s.gpre3 = c.f(Nn * 2 * H); s.gpspr = c.f(Nn * 2 * H);
  s.gx0 = c.f(Nn * 3 * D); s.go1 = c.f(Nn * H); s.go2 = c.f(Nn * H);
  s.gpred = c.f(Nn * D); s.gRglob = c.f(Nn * D * D);
  s.ssend = c.f(Nn * geom_gs(d.num_dims)); s.srecv = c.f(Nn * geom_gr(d.num_dims));
  s.gf = c.f(Nn * D); s.gfz1 = c.f(Nn * 32); s.gfz2 = c.f(Nn * 32); s.gfin = c.f(Nn * (2 * D + 16));
  s.gxin = c.f(Nn * D); s.gvin = c.f(Nn * D);
  s.partial_floats = partial_need(d, flags);
  s.partial = c.f(s.partial_floats);
  s.bytes = c.off;
  return s;
}

// The chained edge-MLP kernel is the default; AETHER_B200_NO_CHAIN=1 runs the two separate GEMMs it replaced (A/B
// measurements only -- the results are bit-identical, tests/test_chain_gpu.py).
bool use_chain() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("AETHER_B200_NO_CHAIN"); v = (e && e[0] == '1') ? 0 : 1; }
  return v == 1;
}

// AETHER_B200_BOX_GATHER=1: the caller asserts that (nearly) every 128-edge tile of the receiver-sorted list has senders
// within 128 consecutive node rows (batched small graphs); the forward then feeds P_s by TMA boxes.  Off by default: it
// measured 8.8 vs 9.1 us per tile (profiles/r02_chain_bench_box.txt) -- the gather variant's cost is the epilogue's
// instruction count, not the sender loads -- and tiles that do not fit take an unprefetched path.
int box_gather_hint() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("AETHER_B200_BOX_GATHER"); v = (e && e[0] == '1') ? 1 : 0; }
  return v;
}

bool dims_ok(const AetherDims* d) {
  return d && (d->num_dims == 2 || d->num_dims == 3) && d->hidden >= 4 && d->hidden % 4 == 0 && d->n_nodes >= 0 &&
         d->n_edges >= 0 && d->n_nodes < ((int64_t)1 << 31) && d->n_edges < ((int64_t)1 << 31);
}

}  // namespace
}  // namespace aether
#include "step_node_tc.inl"
namespace aether {
namespace {

// y = epi(tf(A) W^T + b)
int linear(const NodeTc& nc, const float* A, int64_t lda, int64_t M, int K, const float* W, int64_t ldw, const float* b,
           int N, float* C, int64_t ldc, int tf, const float* mask, const float* res, int64_t ldr, cudaStream_t st) {
  const int rc = linear_tc(nc, A, lda, M, K, W, ldw, b, N, C, ldc, tf, mask, res, ldr, st);
  if (rc != AETHER_E_UNSUPPORTED) return rc;
  GemmArgs g;
  g.A = A; g.lda = lda; g.B = W; g.ldb = ldw; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  g.tf = tf; g.tf_mask = mask; g.bias = b; g.res = res; g.ldr = ldr;
  return gemm_f32(GEMM_NT, g, st);
}

// Row range and phase selection for the receiver-partitioned step.  Default: everything.
//   phase p (0..3) = message-passing layer p+1 on the local edges, its node update on the owned rows and -- for
//   p < 3 -- the node terms of the next layer: P_s[owned] into the exchange buffer (all-gathered by the caller
//   before the next phase), P_r[owned] locally.  Phase 0 starts with field / frames (all nodes: senders need them)
//   and the edge features; phase 3 ends with the out MLP, the rotation back and the Euler update of the owned rows.
// Every bf16 weight image of the step in ONE launch: forward (edge + node), and -- when the backward follows --
// the transposed ones, so that neither pass packs per GEMM (f4: images are built once per step, from the parameters
// the optimizer last wrote).
int pack_step_weights(const AetherDims& d, const float* const* P, const Workspace& w, uint32_t flags, cudaStream_t st) {
  const int H = d.hidden, Fe = num_edge_feat(d.num_dims);
  TcPackJob jobs[TC_PACK_MAX_JOBS];
  int n = 0;
  auto add = [&](const float* W, int ldw, int N, int K, int nv, int kv, int tr, int kmod, void* out) {
    TcPackJob j; j.W = W; j.out = out; j.ldw = ldw; j.N = N; j.K = K; j.n_valid = nv; j.k_valid = kv;
    j.transpose = (int16_t)tr; j.kmod = (int16_t)kmod; jobs[n++] = j;
  };
  const bool save = flags & AETHER_F_SAVE_FOR_BACKWARD;
  add(P[P_L1_M0_W], Fe, H, 64, H, Fe, 0, 32, w.wpA[1]);   // layer 1: the feature columns repeated for the hi | lo halves
  for (int l = 1; l <= 4; ++l) {
    const LayerSlots s = layer_slots(l);
    if (l > 1) add(P[s.m0w] + 2 * H, 3 * H, H, H, H, H, 0, 0, w.wpA[l]);
    add(P[s.m2w], H, H, H, H, H, 0, 0, w.wpB[l]);
    for (int h = 0; h < 2; ++h) {
      add(P[s.u0w] + (size_t)h * H * H, H, H, H, H, H, 0, 0, w.nw3[l][h]);
      add(P[s.u2w] + (size_t)h * H, 2 * H, H, H, H, H, 0, 0, w.nw4[l][h]);
    }
    if (l < 4) {
      const LayerSlots sn = layer_slots(l + 1);
      add(P[sn.m0w], 3 * H, H, H, H, H, 0, 0, w.nws[l]);
      add(P[sn.m0w] + H, 3 * H, H, H, H, H, 0, 0, w.nwr[l]);
    }
    if (save) {
      add(P[s.m2w], H, H, H, H, H, 1, 0, w.wtB[l]);
      if (l > 1) {
        add(P[s.m0w] + 2 * H, 3 * H, H, H, H, H, 1, 0, w.wtA[l]);
        add(P[s.m0w], 3 * H, H, H, H, H, 1, 0, w.nts[l]);            // node terms' data gradients
        add(P[s.m0w] + H, 3 * H, H, H, H, H, 1, 0, w.ntr[l]);
      } else add(P[s.m0w], Fe, 32, H, Fe, H, 1, 0, w.wtA[1]);        // gfeat[E,32] = g1 W1 (fp32 output)
      for (int h = 0; h < 2; ++h) {
        add(P[s.u2w] + (size_t)h * H, 2 * H, H, H, H, H, 1, 0, w.nt4[l][h]);
        add(P[s.u0w] + (size_t)h * H * H, H, H, H, H, H, 1, 0, w.nt3[l][h]);
      }
    }
  }
  return tc_pack_weights_batched(jobs, n, st);
}

struct Part {
  int64_t n0 = 0, n1 = -1;      // owned node rows [n0, n1); n1 < 0: all
  int phase_lo = 0, phase_hi = 3;
  const void* xchg_in = nullptr;   // [>= Nn, H] sender terms P_s read by this phase's edge GEMM (fp32 / bf16)
  void* xchg_out = nullptr;        // ... written (owned rows) for the next phase; null: workspace
};

int forward(const AetherDims& d, const float* const* P, const AetherGraph& G, const float* x, const float* vel,
            const float* q, const float* orig, const float* dmask, float* out, const Workspace& w, uint32_t flags,
            cudaStream_t st, const Part& pt = Part()) {
  const int D = d.num_dims, H = d.hidden, Fe = num_edge_feat(D), ldf = edge_feat_ld(D);
  const int64_t Nn = d.n_nodes, E = d.n_edges;
  const bool bf16 = use_tc(d, flags);
  const int64_t n0 = pt.n0, n1 = pt.n1 < 0 ? Nn : pt.n1, M = n1 - n0;
  NodeTc nc;
  nc.on = bf16; nc.a16 = w.na16; nc.wp = w.nwp;
  // node terms of the factorised first Linear: two planes of [Nn, H], gathered by sender / receiver
  const float* ps32 = pt.xchg_in && !bf16 ? (const float*)pt.xchg_in : w.pspr;
  float* ps32o = pt.xchg_out && !bf16 ? (float*)pt.xchg_out : w.pspr;
  float* pr32 = w.pspr + (size_t)Nn * H;
  const __nv_bfloat16* ps16 = pt.xchg_in && bf16 ? (const __nv_bfloat16*)pt.xchg_in : (const __nv_bfloat16*)w.pspr16;
  __nv_bfloat16* ps16o = pt.xchg_out && bf16 ? (__nv_bfloat16*)pt.xchg_out : (__nv_bfloat16*)w.pspr16;
  __nv_bfloat16* pr16 = (__nv_bfloat16*)w.pspr16 + (size_t)Nn * H;
  if (pt.phase_lo == 0) {
    // neural field at the particles (aether.py:127-134)
    if (flags & AETHER_F_EXTERNAL_FIELD) {
      // DynamicFieldAether: the field comes from the latent field network (dynfield.cu)
      ProfScope pn(PROF_NODE, st);
      AE_CHECK((int)cudaMemcpyAsync(w.f, P[P_F0_W], (size_t)Nn * D * sizeof(float), cudaMemcpyDeviceToDevice, st));
      AE_CHECK(launch_node_frame(D, Nn, vel, w.f, w.R, w.x0, st));
    } else {
    ProfScope pn(PROF_NODE, st);
    AE_CHECK(launch_field_mlp_fwd(D, Nn, x, vel, q, P[P_EMB], P[P_F0_W], P[P_F0_B], P[P_F2_W], P[P_F2_B], P[P_F4_W], P[P_F4_B],
                                  w.fin, w.fz1, w.fz2, w.f, st));
    // local frames and features (aether.py:33-100,177)
    AE_CHECK(launch_node_frame(D, Nn, vel, w.f, w.R, w.x0, st));
    }
    {
      ProfScope ps(PROF_EDGE_FEAT, st);
      if (bf16) AE_CHECK(launch_edge_features_hilo(D, E, G.send, G.recv, G.perm, x, vel, w.f, w.R, w.x0, orig, w.feat16, st));
      else AE_CHECK(launch_edge_features(D, E, G.send, G.recv, G.perm, x, vel, w.f, w.R, w.x0, orig, w.feat, ldf, st));
    }
    if (bf16) {
      ProfScope pn(PROF_NODE, st);
      AE_CHECK(pack_step_weights(d, P, w, flags, st));
    }
  }
  // four message-passing layers (locs.py:187-190, 227-243)
  for (int l = pt.phase_lo + 1; l <= pt.phase_hi + 1; ++l) {
    const LayerSlots s = layer_slots(l);
    const float* xin = w.xl[l - 1];
    {
    ProfScope ps(PROF_EDGE_MLP_FWD, st);
    if (bf16) {
      // pre1 = [silu(pre2_{l-1}) | feat] We^T + b1 (+ P_s[send] + P_r[recv]) ; pre2 = silu(pre1) W2^T + b2
      if (use_chain()) {
        // ONE launch per layer (tc_chain.cu): the intermediate stays in tensor memory and reaches HBM only when the
        // backward needs it
        TcChainArgs c;
        c.M = E; c.wpack1 = w.wpA[l]; c.wpack2 = w.wpB[l]; c.bias1 = P[s.m0b]; c.bias2 = P[s.m2b];
        // layers 2-4: b1 is already inside P_r (bias of the node GEMM that produced it)
        if (l > 1) { c.gat16 = ps16; c.ldg = H; c.idx_a = G.send; c.idx_b = G.recv; c.goff_b = pr16 - ps16; c.bias1 = nullptr; c.gat_rows = Nn; c.box_gather = box_gather_hint(); }
        c.out1 = (flags & AETHER_F_SAVE_FOR_BACKWARD) ? (__nv_bfloat16*)w.bpre1[l] : nullptr; c.ldo1 = H;
        c.out2 = (__nv_bfloat16*)w.bpre2[l]; c.ldo2 = H;
        if (l == 1) AE_CHECK(tc_chain(c, w.feat16, 64, 64, H, TC_TF_NONE, st));
        else AE_CHECK(tc_chain(c, w.bpre2[l - 1], H, H, H, TC_TF_SILU, st));
      } else {
      TcGemmArgs a;
      a.M = E; a.wpack = w.wpA[l]; a.bias = P[s.m0b]; a.out_bf16 = (__nv_bfloat16*)w.bpre1[l]; a.ldo = H;
      if (l > 1) { a.gat16 = ps16; a.ldg = H; a.idx_a = G.send; a.idx_b = G.recv; a.goff_b = pr16 - ps16; }
      if (l == 1) AE_CHECK(tc_gemm(a, w.feat16, 64, 64, H, TC_TF_NONE, st));
      else AE_CHECK(tc_gemm(a, w.bpre2[l - 1], H, H, H, TC_TF_SILU, st));
      TcGemmArgs b;
      b.M = E; b.wpack = w.wpB[l]; b.bias = P[s.m2b]; b.out_bf16 = (__nv_bfloat16*)w.bpre2[l]; b.ldo = H;
      AE_CHECK(tc_gemm(b, w.bpre1[l], H, H, H, TC_TF_SILU, st));
      }
    } else {
      GemmArgs g;
      if (l == 1) {
        g.A = w.feat; g.lda = ldf; g.K = Fe; g.B = P[s.m0w]; g.ldb = Fe; g.tf = TF_NONE;
      } else {
        g.A = w.pre2[l - 1]; g.lda = H; g.K = H; g.B = P[s.m0w] + 2 * H; g.ldb = 3 * H; g.tf = TF_SILU;
        g.gat = ps32; g.ldg = H; g.idx_a = G.send; g.idx_b = G.recv; g.goff_b = pr32 - ps32;
      }
      g.C = w.pre1[l]; g.ldc = H; g.M = E; g.N = H; g.bias = P[s.m0b];
      AE_CHECK(gemm_f32(GEMM_NT, g, st));
      AE_CHECK(linear(nc, w.pre1[l], H, E, H, P[s.m2w], H, P[s.m2b], H, w.pre2[l], H, TF_SILU, nullptr, nullptr, 0, st));
    }
    }
    // y = res(x) + mean_recv(silu(pre2))   (locs.py:236-240), owned rows only
    const float* base = xin + (size_t)n0 * H;
    float* y = w.y[l] + (size_t)n0 * H;
    if (l == 1) {
      ProfScope pn(PROF_NODE, st);
      AE_CHECK(linear(nc, w.x0 + (size_t)n0 * 3 * D, 3 * D, M, 3 * D, P[P_L1_RES_W], 3 * D, P[P_L1_RES_B], H, y, H, TF_NONE,
                      nullptr, nullptr, 0, st));
      base = y;
    }
    const bool node_chain = bf16 && use_chain();
    __nv_bfloat16* y16 = node_chain ? (__nv_bfloat16*)w.y16[l] + (size_t)n0 * H : nullptr;
    {
      ProfScope ps(PROF_SEGMEAN_FWD, st);
      if (bf16) AE_CHECK(launch_segmean_silu_bf16(M, H, G.rowptr + n0, w.bpre2[l], H, base, H, y, H, st, y16, H));
      else AE_CHECK(launch_segmean_silu(M, H, G.rowptr + n0, w.pre2[l], H, base, H, y, H, st));
    }
    // x = y + W4 silu(W3 y + b3) + b4   (locs.py:241)
    ProfScope pn(PROF_NODE, st);
    float* pre3 = w.pre3[l] + (size_t)n0 * 2 * H;
    float* xo = w.xl[l] + (size_t)n0 * H;
    if (node_chain) {
      // the update MLP is GEMM -> SiLU -> GEMM like the edge MLP: two launches of the chained kernel, one per half of the
      // 2H-wide hidden layer (x = y + b4 + silu(pre3_a) W4a^T + silu(pre3_b) W4b^T), fp32 residual output, bf16 copies of
      // pre3 (backward) and x (operand of the next layer's node terms) written by the epilogues
      __nv_bfloat16* p316 = (__nv_bfloat16*)w.pre316[l] + (size_t)n0 * 2 * H;
      __nv_bfloat16* x16 = (__nv_bfloat16*)w.x16[l] + (size_t)n0 * H;
      for (int h = 0; h < 2; ++h) {
        TcChainArgs c;
        c.M = M; c.wpack1 = w.nw3[l][h]; c.wpack2 = w.nw4[l][h]; c.bias1 = P[s.u0b] + h * H; c.bias2 = h == 0 ? P[s.u2b] : nullptr;
        c.out1 = (flags & AETHER_F_SAVE_FOR_BACKWARD) ? p316 + h * H : nullptr; c.ldo1 = 2 * H;
        c.out2_f32 = xo; c.ldo2f = H; c.res = h == 0 ? y : xo; c.ldr = H;
        if (h == 1) { c.copy16 = x16; c.ldc16 = H; }
        AE_CHECK(tc_chain(c, y16, H, H, H, TC_TF_NONE, st));
      }
      if (l < 4) {
        const LayerSlots sn = layer_slots(l + 1);
        TcGemmArgs a;
        a.M = M; a.wpack = w.nws[l]; a.out_bf16 = ps16o + (size_t)n0 * H; a.ldo = H;
        AE_CHECK(tc_gemm(a, x16, H, H, H, TC_TF_NONE, st));
        TcGemmArgs b;
        b.M = M; b.wpack = w.nwr[l]; b.out_bf16 = pr16 + (size_t)n0 * H; b.ldo = H;
        b.bias = P[sn.m0b];            // P_r + b1: the chained edge kernel of layer l+1 then adds no bias of its own
        AE_CHECK(tc_gemm(b, x16, H, H, H, TC_TF_NONE, st));
      }
      continue;
    }
    AE_CHECK(linear(nc, y, H, M, H, P[s.u0w], H, P[s.u0b], 2 * H, pre3, 2 * H, TF_NONE, nullptr, nullptr, 0, st));
    AE_CHECK(linear(nc, pre3, 2 * H, M, 2 * H, P[s.u2w], 2 * H, P[s.u2b], H, xo, H, TF_SILU, nullptr, y, H, st));
    if (l < 4) {
      // node-factorised first Linear of the next layer: [x_s | x_r | e] W^T = x_s Ws^T + x_r Wr^T + e We^T
      const LayerSlots sn = layer_slots(l + 1);
      if (bf16) {
        AE_CHECK(linear_tc_bf16out(nc, xo, H, M, H, P[sn.m0w], 3 * H, H, ps16o + (size_t)n0 * H, H, st));
        AE_CHECK(linear_tc_bf16out(nc, xo, H, M, H, P[sn.m0w] + H, 3 * H, H, pr16 + (size_t)n0 * H, H, st));
      } else {
        AE_CHECK(linear(nc, xo, H, M, H, P[sn.m0w], 3 * H, nullptr, H, ps32o + (size_t)n0 * H, H, TF_NONE, nullptr, nullptr, 0, st));
        AE_CHECK(linear(nc, xo, H, M, H, P[sn.m0w] + H, 3 * H, nullptr, H, pr32 + (size_t)n0 * H, H, TF_NONE, nullptr, nullptr, 0, st));
      }
    }
  }
  if (pt.phase_hi < 3) return 0;
  // out MLP (locs.py:160-168,193), globalize and Euler update (local_to_global.py:12-13, aether.py:185)
  ProfScope pn(PROF_NODE, st);
  const float* m1 = dmask ? dmask + (size_t)n0 * H : nullptr;
  const float* m2 = dmask ? dmask + (size_t)Nn * H + (size_t)n0 * H : nullptr;
  float* o1 = w.o1 + (size_t)n0 * H;
  float* o2 = w.o2 + (size_t)n0 * H;
  float* pred = w.pred + (size_t)n0 * D;
  AE_CHECK(linear(nc, w.xl[4] + (size_t)n0 * H, H, M, H, P[P_OUT0_W], H, P[P_OUT0_B], H, o1, H, TF_NONE, nullptr, nullptr, 0, st));
  AE_CHECK(linear(nc, o1, H, M, H, P[P_OUT3_W], H, P[P_OUT3_B], H, o2, H, m1 ? TF_SILU_MASK : TF_SILU, m1, nullptr, 0, st));
  AE_CHECK(linear(nc, o2, H, M, H, P[P_OUT6_W], H, P[P_OUT6_B], D, pred, D, m2 ? TF_SILU_MASK : TF_SILU, m2, nullptr, 0, st));
  AE_CHECK(launch_globalize(D, M, x + (size_t)n0 * D, w.R + (size_t)n0 * D * D, pred, out + (size_t)n0 * D, st));
  return 0;
}

}  // namespace
}  // namespace aether

#include "step_backward.inl"
#include "step_abi.inl"
