/* Plain-C caller of the aether_b200 ABI: buffer sizes for one configuration.
 *
 *   gcc -std=c99 -I include examples/abi_sizes.c -L aether_b200/lib -laether_b200 -o abi_sizes
 *   LD_LIBRARY_PATH=aether_b200/lib ./abi_sizes 3 256 102400 10137600
 *
 * Only the size queries are called, so this runs without a GPU; a real caller then cudaMallocs the workspace and
 * scratch it reports and calls aether_graph_prepare / aether_step_forward / aether_step_backward on its stream.
 */
#include <stdio.h>
#include <stdlib.h>

#include "aether_b200.h"

int main(int argc, char** argv) {
  AetherDims d;
  d.num_dims = argc > 1 ? atoi(argv[1]) : 3;
  d.hidden = argc > 2 ? atoi(argv[2]) : 64;
  d.n_nodes = argc > 3 ? atoll(argv[3]) : 640;
  d.n_edges = argc > 4 ? atoll(argv[4]) : 2560;
  if (aether_abi_version() != AETHER_B200_ABI_VERSION) {
    fprintf(stderr, "ABI version mismatch\n");
    return 2;
  }
  const size_t infer = aether_workspace_bytes(&d, 0);
  const size_t train = aether_workspace_bytes(&d, AETHER_F_SAVE_FOR_BACKWARD);
  const size_t train16 = aether_workspace_bytes(&d, AETHER_F_SAVE_FOR_BACKWARD | AETHER_F_BF16_EDGE_MLP);
  const size_t scratch = aether_backward_scratch_bytes(&d, AETHER_F_SAVE_FOR_BACKWARD);
  const size_t graph = aether_graph_scratch_bytes(d.n_nodes, d.n_edges);
  if (infer == 0) {
    fprintf(stderr, "unsupported dimensions: %s\n", aether_error_string(AETHER_E_BADARG));
    return 1;
  }
  printf("abi %d  D=%d H=%d Nn=%lld E=%lld\n", aether_abi_version(), d.num_dims, d.hidden, (long long)d.n_nodes,
         (long long)d.n_edges);
  printf("workspace inference %zu  training fp32 %zu  training bf16 %zu  backward scratch %zu  graph scratch %zu\n", infer,
         train, train16, scratch, graph);
  /* the seq2seq decoder path at the shape of scripts/electrostatic_field_aether.sh: 5 particles, batch 64, decoder and
   * encoder hidden 512, LSTM 128, prior head 256, 2 edge types */
  {
    AetherSeqDims s;
    s.num_dims = d.num_dims == 3 ? 3 : 2;
    s.num_vars = 5;
    s.hidden = 512;
    s.n_types = 2;
    s.skip_first = 0;
    s.field_hidden = 512;
    s.rnn_hidden = 128;
    s.prior_hidden = 256;
    s.flags = 0;
    s.batch = 64;
    printf("seq2seq decoder + encoder workspace %zu\n", aether_seq_workspace_bytes(&s));
  }
  return 0;
}
