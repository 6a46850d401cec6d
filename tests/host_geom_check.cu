// CPU harness around the __host__ __device__ geometry bodies of aether_b200/csrc/geom.cu and
// geom_bwd.cu: the not-gpu tests call these loops to check the forward features and the
// hand-derived backward against the oracle / autograd without a GPU.  Test infrastructure only.
#include "../aether_b200/csrc/geom.cu"
#include "../aether_b200/csrc/geom_bwd.cu"

namespace aether { std::atomic<unsigned long long> g_launch_count{0}; }
using namespace aether;

template <int D>
static void fwd(int64_t Nn, int64_t E, const int32_t* send, const int32_t* recv, const int32_t* perm, const float* x,
                const float* vel, const float* f, const float* orig, float* R, float* x0, float* feat) {
  for (int64_t n = 0; n < Nn; ++n) node_frame_body<D>(n, vel, f, R, x0);
  for (int64_t e = 0; e < E; ++e) edge_features_body<D>(e, send, recv, perm, x, vel, f, R, x0, orig, feat, edge_feat_ld(D));
}

template <int D>
static void bwd(int64_t Nn, int64_t E, const int32_t* send, const int32_t* recv, const float* x, const float* vel,
                const float* f, const float* R, const float* pred, const float* gfeat, const float* gx0,
                const float* gout, float* gsend, float* grecv, float* ssend, float* srecv, float* gRglob, float* gpred,
                float* gf, float* gx, float* gvel) {
  const int GS = geom_gs(D);
  for (int64_t n = 0; n < Nn; ++n) globalize_bwd_body<D>(n, R, pred, gout, gpred, gRglob);
  for (int64_t e = 0; e < E; ++e) edge_features_bwd_body<D>(e, send, recv, x, vel, f, R, gfeat, edge_feat_ld(D), gsend, grecv);
  for (int64_t i = 0; i < Nn * GS; ++i) ssend[i] = srecv[i] = 0.f;
  for (int64_t e = 0; e < E; ++e)
    for (int c = 0; c < GS; ++c) {
      ssend[(int64_t)send[e] * GS + c] += gsend[e * GS + c];
      srecv[(int64_t)recv[e] * GS + c] += grecv[e * GS + c];
    }
  for (int64_t n = 0; n < Nn; ++n) node_frame_bwd_body<D>(n, vel, f, R, gx0, ssend, srecv, gRglob, gout, gf, gx, gvel);
}

extern "C" void hostcheck_geom_forward(int D, int64_t Nn, int64_t E, const int32_t* send, const int32_t* recv,
                                       const int32_t* perm, const float* x, const float* vel, const float* f,
                                       const float* orig, float* R, float* x0, float* feat) {
  if (D == 2) fwd<2>(Nn, E, send, recv, perm, x, vel, f, orig, R, x0, feat);
  else fwd<3>(Nn, E, send, recv, perm, x, vel, f, orig, R, x0, feat);
}

// Scratch arrays are caller-allocated: gsend/grecv [E,GS], ssend/srecv [Nn,GS], gRglob [Nn,D*D].
extern "C" void hostcheck_geom_backward(int D, int64_t Nn, int64_t E, const int32_t* send, const int32_t* recv,
                                        const float* x, const float* vel, const float* f, const float* R,
                                        const float* pred, const float* gfeat, const float* gx0, const float* gout,
                                        float* gsend, float* grecv, float* ssend, float* srecv, float* gRglob,
                                        float* gpred, float* gf, float* gx, float* gvel) {
  if (D == 2) bwd<2>(Nn, E, send, recv, x, vel, f, R, pred, gfeat, gx0, gout, gsend, grecv, ssend, srecv, gRglob, gpred, gf, gx, gvel);
  else bwd<3>(Nn, E, send, recv, x, vel, f, R, pred, gfeat, gx0, gout, gsend, grecv, ssend, srecv, gRglob, gpred, gf, gx, gvel);
}
