// placeholder until the tcgen05 kernel lands (next commit)
#include "ucn_common.cuh"
int ucn_mlp_forward_tc(const ucn_weights *, const float *, const float *, const float *, int, int, float *, void *) {
  return ucn_set_error(UCN_E_UNSUPPORTED, "ucn_mlp_forward: BF16_TC mode not built yet");
}
