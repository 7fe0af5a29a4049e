// Multi-GPU PS step (push / apply+pull over NVLink peer memory).
#include "ps_common.h"

using accel::fail;

extern "C" {

int accel_shard_push(accel_ctx *, const accel_peers *, const void *const *, int, int,
                     accel_stream_t) {
  return fail(ACCEL_ERR_STATE, "accel_shard_push: not built yet");
}

int accel_shard_apply_pull(accel_ctx *, const accel_peers *, int, float *, float *, int,
                           const accel_hparams *, const float *, accel_stream_t) {
  return fail(ACCEL_ERR_STATE, "accel_shard_apply_pull: not built yet");
}

int accel_shard_wait_pull(accel_ctx *, const accel_peers *, accel_stream_t) {
  return fail(ACCEL_ERR_STATE, "accel_shard_wait_pull: not built yet");
}

}  // extern "C"
