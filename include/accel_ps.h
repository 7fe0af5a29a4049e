/*
 * accel_ps.h — C ABI of the B200-native parameter-server (PS) step.
 *
 * This is the drop-in boundary for ONE path of gaocegege/kubeflow-controller's
 * TFJobs: on each PS task, aggregate the gradients pushed by N workers for the
 * task's variable shard and apply the optimizer update (SGD / Momentum / Adam).
 * The reference has no FFI of its own (it is a Go control plane; the arithmetic
 * lives in TensorFlow 1.4.0, pinned by image tag examples/tfjob/dist.yml:13,32),
 * so every entry point below cites the reference interface or call site whose
 * behaviour it reproduces. The Go package pkg/accel binds exactly these symbols
 * through cgo.
 *
 * Conventions
 *   - plain C types only; no exceptions cross the boundary;
 *   - return value 0 = ACCEL_OK, negative = accel_status error code; the
 *     message of the last error on the calling thread is accel_last_error();
 *   - every `*_dev` / device pointer is CALLER-OWNED device memory, never freed
 *     or retained past the call; state/gradient pointers must be 16-byte
 *     aligned (32-byte for ACCEL_KERNEL_LDG256);
 *   - `stream` is a cudaStream_t passed as an opaque pointer (NULL = the legacy
 *     default stream); kernels are enqueued asynchronously on it;
 *   - PS task_index i (pkg/tensorflow/distributed.go:151 "--task_index=%d")
 *     <-> GPU ordinal gpu_ordinals[i] <-> collective rank i <-> shard i;
 *   - calls on the same task must be serialised by the caller; different tasks
 *     may be driven from different host threads (each call sets its device);
 *   - there is NO CPU fallback: without a usable CUDA device every compute
 *     entry point fails with ACCEL_ERR_CUDA.
 */
#ifndef ACCEL_PS_H_
#define ACCEL_PS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ACCEL_PS_VERSION 10200 /* major*10000 + minor*100 + patch */
#define ACCEL_MAX_WORKERS 64   /* gradients summed by one launch */
#define ACCEL_MAX_PS 16        /* PS tasks (= GPUs of one box) per context */

typedef struct accel_ctx accel_ctx;
typedef void *accel_stream_t; /* cudaStream_t */

typedef enum accel_status {
  ACCEL_OK = 0,
  ACCEL_ERR_INVALID = -1, /* bad argument (NULL, out of range, unknown enum) */
  ACCEL_ERR_ALIGN = -2,   /* pointer not aligned for the vector path */
  ACCEL_ERR_CUDA = -3,    /* CUDA runtime error (message has the detail) */
  ACCEL_ERR_NOMEM = -4,   /* allocation failed */
  ACCEL_ERR_BUFSIZE = -5, /* caller's output buffer too small */
  ACCEL_ERR_STATE = -6    /* call not valid in the context's current state */
} accel_status;

typedef enum accel_dtype { ACCEL_F32 = 0, ACCEL_BF16 = 1 } accel_dtype;

/* How the N worker gradients become one gradient.
 * SUM  : g = ((g0 + g1) + g2) + ...  in worker-index order, fp32.
 * MEAN : the SUM divided by float(N) — what TF-1.4's ConditionalAccumulator
 *        take_grad returns under SyncReplicasOptimizer
 *        (examples/workdir/mnist_replica.py:172-182). */
typedef enum accel_aggregate { ACCEL_AGG_SUM = 0, ACCEL_AGG_MEAN = 1 } accel_aggregate;

typedef enum accel_optimizer {
  ACCEL_OPT_SGD = 0,      /* tf.train.GradientDescentOptimizer, examples/workdir/mnist_softmax.py:59 */
  ACCEL_OPT_MOMENTUM = 1, /* tf.train.MomentumOptimizer (BASELINE config 3) */
  ACCEL_OPT_ADAM = 2      /* tf.train.AdamOptimizer, examples/workdir/mnist_replica.py:170 */
} accel_optimizer;

/* Kernel variant: an explicit choice of the caller (benchmarks sweep it), never
 * a dispatch on hardware. AUTO = the variant measured fastest on B200. */
typedef enum accel_kernel {
  ACCEL_KERNEL_AUTO = 0,
  ACCEL_KERNEL_LDG128 = 1, /* 128-bit streaming loads, grid-stride */
  ACCEL_KERNEL_LDG256 = 2, /* 256-bit streaming loads (sm_100 LDG.256) */
  ACCEL_KERNEL_TMA = 3     /* 1-D bulk-async (TMA) staged tiles + mbarrier ring */
} accel_kernel;

/* Hyper-parameters, by value. Scalars are fp32 exactly as TF's CPU kernels take
 * them (TF-1.4 training_ops.cc ApplyGradientDescent / ApplyMomentum /
 * ApplyAdamNonCuda; call sites mnist_replica.py:170, mnist_softmax.py:59). */
typedef struct accel_hparams {
  float lr;
  float momentum;     /* Momentum */
  float beta1;        /* Adam */
  float beta2;        /* Adam */
  float epsilon;      /* Adam */
  float beta1_power;  /* Adam: beta1^t, initial value beta1 (TF adam.py _create_slots) */
  float beta2_power;  /* Adam: beta2^t, initial value beta2 */
  float clip_norm;    /* > 0 enables tf.clip_by_global_norm on the aggregated gradient */
  int32_t use_nesterov; /* Momentum and Adam */
  int32_t aggregate;    /* accel_aggregate */
} accel_hparams;

typedef struct accel_shard {
  uint64_t offset; /* first element of the shard in the padded flat arena */
  uint64_t count;  /* elements in the shard (padding included) */
} accel_shard;

/* ---- library ------------------------------------------------------------ */

/* Library version (ACCEL_PS_VERSION). No reference counterpart: the reference
 * reports its own version via version/version.go:3-6. */
int accel_version(void);

/* Message of the last failing call on this thread ("" if none). Mirrors Go's
 * `error` return convention (pkg/controller/controller.go:237-240). */
const char *accel_last_error(void);

/* ---- context: PS task -> GPU map ---------------------------------------- */

/* Create a context for n_ps parameter-server tasks. gpu_ordinals[i] is the CUDA
 * device of PS task i; NULL means the identity map. n_ps is the `replicas` of
 * the PS TFReplicaSpec (vendor/github.com/caicloud/kubeflow-clientset/apis/
 * kubeflow/v1alpha1/types.go:58-79, read at pkg/tensorflow/distributed.go:140).
 * No device is touched until a task is first used. */
int accel_ctx_create(const int *gpu_ordinals, int n_ps, accel_ctx **out);
int accel_ctx_destroy(accel_ctx *ctx);

/* GPU ordinal serving PS task `task` (the "--task_index" emitted at
 * pkg/tensorflow/distributed.go:151); negative accel_status on error. */
int accel_ctx_gpu_of(const accel_ctx *ctx, int task);

/* Select kernel variant and launch shape for subsequent applies. threads = CTA
 * size (multiple of 32, <= 1024), ctas_per_sm = resident CTAs per SM of the
 * persistent grid, unroll = independent vector loads per stream per thread.
 * Any argument 0 keeps the built-in default. Benchmark/tuning knob; no
 * reference counterpart. */
int accel_ctx_set_kernel(accel_ctx *ctx, int kernel, int threads, int ctas_per_sm, int unroll);

/* ---- variable -> PS task sharding (SURVEY §8 a7, §8 e) ------------------ */

/* Flat mode: pad n_elems up to a multiple of n_ps*align_elems and split into
 * n_ps equal contiguous shards (reduce-scatter ownership). out[n_ps]. */
int accel_plan_flat(uint64_t n_elems, int n_ps, uint64_t align_elems, accel_shard *out,
                    uint64_t *padded_elems);

/* TF-faithful mode: tf.train.replica_device_setter's default round-robin —
 * variable k (creation order) goes to PS task k mod n_ps
 * (examples/workdir/mnist_replica.py:137-158). out_task[n_vars] receives the
 * task, out_offset[n_vars] the variable's first element inside its task's
 * bucket (each variable start rounded up to align_elems), bucket_elems[n_ps]
 * the packed size of each bucket. `first_task` is the round-robin counter at
 * the first of these variables (non-trainable variables created earlier, e.g.
 * global_step at mnist_replica.py:142, advance it). */
int accel_plan_round_robin(const uint64_t *var_elems, int n_vars, int n_ps, int first_task,
                           uint64_t align_elems, int *out_task, uint64_t *out_offset,
                           uint64_t *bucket_elems);

/* ---- the hot path: fused N-way gradient aggregation + optimizer apply ---- */

/* var -= lr * G, where G = aggregate(grads[0..n_workers-1]) (then clipped when
 * hp->clip_norm > 0). One pass over HBM: n_workers*sizeof(g) + 8 bytes/element.
 * Replaces TF-1.4 ConditionalAccumulator + ApplyGradientDescent on the PS CPU
 * (reference call site examples/workdir/mnist_softmax.py:59; SURVEY §8 a8,a9).
 * grads: host array of n_workers device pointers, each `count` elements of
 * grad_dtype. grad_sumsq_dev: device pointer to sum(G^2) over the WHOLE model
 * (all shards) for clipping; NULL with clip_norm > 0 makes the call compute it
 * over [0,count) itself with accel_grad_sumsq into context scratch. */
int accel_apply_sgd(accel_ctx *ctx, int task, float *var, const void *const *grads,
                    int n_workers, int grad_dtype, uint64_t count, const accel_hparams *hp,
                    const float *grad_sumsq_dev, accel_stream_t stream);

/* accum = accum*momentum + G; var -= accum*lr
 * (nesterov: var -= G*lr + accum*momentum*lr). n_workers*sizeof(g) + 16 B/elem.
 * Replaces TF-1.4 ApplyMomentum. */
int accel_apply_momentum(accel_ctx *ctx, int task, float *var, float *accum,
                         const void *const *grads, int n_workers, int grad_dtype,
                         uint64_t count, const accel_hparams *hp, const float *grad_sumsq_dev,
                         accel_stream_t stream);

/* alpha = lr*sqrt(1-beta2_power)/(1-beta1_power);
 * m += (G-m)*(1-beta1); v += (G*G-v)*(1-beta2); var -= (m*alpha)/(sqrt(v)+epsilon)
 * (nesterov: var -= ((G*(1-beta1) + beta1*m)*alpha)/(sqrt(v)+epsilon)).
 * n_workers*sizeof(g) + 24 B/elem. Replaces TF-1.4 ApplyAdam
 * (examples/workdir/mnist_replica.py:170). The caller advances
 * the powers after the step with accel_adam_advance (TF does so in Python). */
int accel_apply_adam(accel_ctx *ctx, int task, float *var, float *m, float *v,
                     const void *const *grads, int n_workers, int grad_dtype, uint64_t count,
                     const accel_hparams *hp, const float *grad_sumsq_dev,
                     accel_stream_t stream);

/* beta1_power *= beta1; beta2_power *= beta2 — TF-1.4 AdamOptimizer._finish,
 * run once per step after all ApplyAdam ops (mnist_replica.py:170,184). */
int accel_adam_advance(accel_hparams *hp);

/* Optimizer state that advances ON THE DEVICE (SURVEY §8 f, rank 2): a 16-byte,
 * 16-byte-aligned, caller-owned device block {float beta1_power, float
 * beta2_power, uint64 global_step}. accel_opt_state_init fills it from hp's
 * powers; accel_apply_dev_state is accel_apply_sgd/_momentum/_adam with Adam's
 * alpha computed in the kernel from the device powers (hp's powers are
 * ignored; same fp32 operations as the host form, so results are identical)
 * followed on the same stream by TF's finish step — powers *= betas (Adam),
 * global_step += 1 (the increment of mnist_replica.py:142,184). Nothing in a
 * step depends on host-side mutable state, so a step can be recorded once and
 * replayed. slot0/slot1 are ignored where the optimizer has no such slot. */
int accel_opt_state_init(void *dev_state, const accel_hparams *hp, uint64_t global_step,
                         accel_stream_t stream);
/* TF's finish step on its own: powers *= betas (when is_adam), global_step += 1.
 * accel_apply_dev_state runs it after EVERY call — correct when a task's
 * variables live in one flat arena (one call per step). A task that applies
 * several variables per step (accel_plan_round_robin buckets applied one by
 * one) uses accel_apply_dev_state_ex with advance = 0 and calls this once. */
int accel_opt_state_advance(void *dev_state, const accel_hparams *hp, int is_adam,
                            accel_stream_t stream);
int accel_apply_dev_state(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                          float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                          uint64_t count, const accel_hparams *hp, const float *grad_sumsq_dev,
                          void *dev_state, accel_stream_t stream);
/* accel_apply_dev_state with the finish step under the caller's control:
 * advance = 0 applies only (call accel_opt_state_advance once after the LAST
 * variable of the step, as TF-1.4's AdamOptimizer._finish does); advance != 0
 * is accel_apply_dev_state. */
int accel_apply_dev_state_ex(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                             float *slot1, const void *const *grads, int n_workers,
                             int grad_dtype, uint64_t count, const accel_hparams *hp,
                             const float *grad_sumsq_dev, void *dev_state, int advance,
                             accel_stream_t stream);

/* out_sumsq_dev[0] = sum over [0,count) of G^2, G = aggregate(grads), for
 * tf.clip_by_global_norm. Deterministic for a given launch
 * shape: per-thread fp32, warp-shuffle + shared-memory tree per CTA, CTA
 * partials to context scratch, summed in index order in fp64 by the last CTA.
 * Uses scratch global memory owned by the context (allocated at first use on
 * the task's device). */
int accel_grad_sumsq(accel_ctx *ctx, int task, const void *const *grads, int n_workers,
                     int grad_dtype, uint64_t count, int aggregate, float *out_sumsq_dev,
                     accel_stream_t stream);

/* accel_grad_sumsq for callers that must not allocate device memory. It has no
 * counterpart in the reference. Who needs it: a PS process living inside a
 * framework that has already handed all of HBM to its own arena at start-up (a
 * later cudaMalloc by this library, such as the scratch accel_ctx allocates on
 * a task's first use, would simply fail there), and callers on a step loop
 * that cannot afford the device-wide synchronisation a cudaMalloc/cudaFree
 * implies. This entry point takes no context, runs on the CURRENT device, and
 * calls no cudaMalloc, cudaFree or other allocating routine: its scratch (one
 * fp32 partial per CTA and a ticket word, 16.25 KiB) is a static device array
 * inside the library, present on every GPU that loads the library until the
 * process exits.
 *   - Limit: count <= ACCEL_SUMSQ_STATIC_MAX_ELEMS (2^28 elements per worker
 *     gradient, i.e. 4096 CTAs of 65536 elements); larger counts are refused
 *     with ACCEL_ERR_INVALID — split the range and add the results, or use
 *     accel_grad_sumsq.
 *   - The first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress, once per device); make that first call during
 *     set-up, before any stream capture begins.
 *   - All calls on a device share the one array. The ticket word is cleared on
 *     the stream at the start of every call; calls whose kernels could overlap
 *     on the GPU (different streams, or different host threads without
 *     ordering) race on it — serialise them.
 *   - Result: same definition as accel_grad_sumsq, same tolerance (the CTA
 *     count differs, so the last bits of the fp32 sum may differ); for a given
 *     count it is reproducible and does not depend on the GPU's SM count. */
#define ACCEL_SUMSQ_STATIC_MAX_ELEMS 268435456ull
int accel_grad_sumsq_static(const void *const *grads, int n_workers, int grad_dtype,
                            uint64_t count, int aggregate, float *out_sumsq_dev,
                            accel_stream_t stream);

/* ---- one PS step with HOST buffers (the reference-facing call) ----------- */

/* What a PS replica does per step when the workers' gradients arrive in host
 * memory (the reference moves them by gRPC into PS host memory,
 * examples/workdir/mnist_replica.py:251-264): copy the n_workers host gradient
 * buffers to the GPU in chunks, run the fused aggregate+apply on each chunk as
 * it lands, and copy the updated variables back to host_var_out (the workers'
 * pull; may be NULL to skip). var/slot0/slot1 are DEVICE resident PS state
 * (slot0 = accum or m, slot1 = v; NULL when unused). Copies and kernels are
 * pipelined on context-owned streams; the call returns after the step is
 * complete. Host buffers should be page-locked (accel_host_register). Global-
 * norm clipping is not available on this entry point (it needs the whole
 * gradient before the first chunk can be applied): clip_norm must be 0. */
int accel_ps_step_host(accel_ctx *ctx, int task, int optimizer, float *var, float *slot0,
                       float *slot1, const void *const *host_grads, int n_workers,
                       int grad_dtype, uint64_t count, const accel_hparams *hp,
                       float *host_var_out);

/* Page-lock / unlock a host buffer so the step above copies at full PCIe rate. */
int accel_host_register(void *host_ptr, size_t bytes);
int accel_host_unregister(void *host_ptr);

/* ---- placement: the contract pkg/tensorflow emits ------------------------ */

typedef enum accel_replica_type {
  ACCEL_REPLICA_PS = 0,
  ACCEL_REPLICA_WORKER = 1,
  ACCEL_REPLICA_LOCAL = 2 /* only meaningful to accel_tfjob_is_local_job */
} accel_replica_type;

/* Action types in the order of pkg/tensorflow/types.go:28-34. */
typedef enum accel_action {
  ACCEL_ACTION_ADD_PS = 0,
  ACCEL_ACTION_ADD_PS_SERVICE = 1,
  ACCEL_ACTION_ADD_WORKER = 2,
  ACCEL_ACTION_ADD_WORKER_SERVICE = 3,
  ACCEL_ACTION_DELETE = 4,
  ACCEL_ACTION_NOTHING = 5
} accel_action;

typedef struct accel_event {
  int32_t action; /* accel_action */
  int32_t number;
} accel_event;

/* Desired-minus-actual replica counts -> ordered events, exactly as
 * DistributedJob.Action (pkg/tensorflow/distributed.go:56-114). out must hold
 * 4 events; *n_out receives the count; *composed is set to 1 when the
 * reference would call compose() (new RuntimeID), else 0. */
int accel_tfjob_action(int worker_replicas, int ps_replicas, int succeeded_workers,
                       int worker_services, int ps_services, int active_workers, int active_ps,
                       accel_event *out, int *n_out, int *composed);

/* checker.IsLocalJob (pkg/checker/checker.go:8-14): a job is local iff its FIRST
 * replica spec has type Local. replica_types[n] are accel_replica_type values in
 * spec order. Returns 1 / 0, or a negative status. */
int accel_tfjob_is_local_job(const int *replica_types, int n);

/* LocalJob.Action (pkg/tensorflow/local.go:50-74) for a job whose first spec has type
 * first_spec_type, with NewLocalJob's pod selection (:34-48: the pod is kept only when there is
 * exactly one active pod) and compose()/labels() (:78-90) when a worker must be added: the new
 * RuntimeID is `suffix` (5 characters, as generateRuntimeID draws them) and the template labels
 * become {kubeflow.caicloud.io, job_type, runtime_id, tf_job_name}. *out_event receives the
 * single event; *composed 1 when compose() ran; buf receives "runtime_id<TAB><id>\n" and then
 * the template's labels as "key=value\n" lines sorted by key (none unless compose() ran).
 * Returns bytes written excluding the NUL, or a negative status. */
int accel_tfjob_local_action(const char *job_name, const char *runtime_id_in, int first_spec_type,
                             int active_pods, int succeeded, const char *suffix,
                             accel_event *out_event, int *composed, char *buf, size_t buflen);

/* One whole sync of a distributed TFJob, as manageTFJob runs it
 * (pkg/controller/controller.go:369-425): NewDistributedJob, Action(), then the
 * events in list order — GetService(type, i) per service, GetSpec(type, i) per
 * pod, indices restarting at 0 for however many are missing — with the random
 * 5-character tails taken from `suffixes` in the order the reference draws
 * them (compose()'s RuntimeID first when it runs, then one per Service).
 * runtime_id_in is the job's stored Spec.RuntimeID (used when compose() does
 * not run). The result is text, one record per line, fields TAB-separated:
 *   runtime_id <TAB> <id after the sync>
 *   event <TAB> <ActionType> <TAB> <number>
 *   service <TAB> <generated name> <TAB> <k=v,k=v,... labels sorted by key>
 *   pod <TAB> <PS|Worker> <TAB> <index> <TAB> <labels> <TAB> <arg0> <TAB> <arg1> <TAB> <arg2> <TAB> <arg3>
 *   template <TAB> <PS|Worker> <TAB> <labels left in the shared template> <TAB> <args left in it>
 * (the `template` lines show the in-place mutation of GetSpec, distributed.go:118-124).
 * Returns bytes written excluding the NUL, or a negative status. */
int accel_tfjob_sync_plan(const char *job_name, const char *runtime_id_in, int worker_replicas,
                          int ps_replicas, int succeeded_workers, int worker_services,
                          int ps_services, int active_workers, int active_ps,
                          const char *const *suffixes, int n_suffixes, char *buf, size_t buflen);

/* "<job>-<type lower-cased>-<index>" — DistributedJob.getServiceName
 * (pkg/tensorflow/distributed.go:161-163). Returns the length written
 * (excluding NUL) or a negative accel_status. */
int accel_tfjob_service_key(const char *job_name, int replica_type, int index, char *buf,
                            size_t buflen);

/* The four argv strings of generateTFClusterSpec
 * (pkg/tensorflow/distributed.go:127-159), NUL-separated in buf in the order
 * --worker_hosts, --ps_hosts, --job_name, --task_index. worker_services[W] /
 * ps_services[P] are the generated Service names (an entry may be "" — the
 * reference then emits ":2222", distributed.go:131,142 with an empty map).
 * Returns total bytes written including the four NULs, or negative status. */
int accel_tfjob_cluster_spec(int worker_replicas, int ps_replicas,
                             const char *const *worker_services, const char *const *ps_services,
                             int replica_type, int index, char *buf, size_t buflen);

/* Pod/Service label set of getLabels (+ "index" when index >= 0):
 * distributed.go:121,165-167,221-228. Written as "key=value\n" lines sorted by
 * key (Go maps are unordered). Returns bytes written excluding NUL. */
int accel_tfjob_labels(const char *job_name, int replica_type, const char *runtime_id, int index,
                       char *buf, size_t buflen);

/* Parse the four flags a replica receives back into (job type, task index,
 * host counts) — what examples/workdir/mnist_replica.py:107-115 does before
 * tf.train.Server — and return the GPU ordinal for a PS task through ctx. */
int accel_tfjob_parse_argv(const char *const *argv, int argc, int *replica_type, int *task_index,
                           int *n_workers, int *n_ps);

/* ---- multi-GPU: PS shards on the GPUs of one NVSwitch box (SURVEY §8 e) -- */

#define ACCEL_IPC_HANDLE_BYTES 64
typedef struct accel_ipc_handle {
  unsigned char bytes[ACCEL_IPC_HANDLE_BYTES];
} accel_ipc_handle;

/* Export a device allocation for mapping by the other PS processes of the box,
 * and map / unmap a peer's. (One process per GPU; handles travel over the
 * rendezvous the launcher already has.) Replaces the reference's gRPC transport
 * between pods (pkg/tensorflow/distributed.go:26-29,178-185, port 2222). */
int accel_ipc_export(void *dev_ptr, accel_ipc_handle *out);
int accel_ipc_open(const accel_ipc_handle *h, void **out_dev_ptr);
int accel_ipc_close(void *dev_ptr);

/* Device memory that can be exported with accel_ipc_export: one whole allocation
 * per buffer (a sub-range of a pooled allocator's block cannot be exported on
 * its own), zero-filled. The caller owns it and frees it with accel_dev_free.
 * No reference counterpart. */
int accel_dev_alloc(size_t bytes, void **out_dev_ptr);
int accel_dev_free(void *dev_ptr);

/* Peer views of one rank's multi-GPU step. All arrays have `world` entries
 * indexed by PS task / rank; entry [rank] is this rank's own local buffer.
 *   stage[r]  : fp32 [world][shard] receive slots on rank r — slot s holds the
 *               partial gradient sum that rank s pushed for shard r;
 *   var[r]    : fp32 [world*shard] full variable replica on rank r (the
 *               workers' pull target);
 *   flags[r]  : uint32 [ACCEL_SHARD_FLAG_WORDS] barrier words on rank r, zeroed
 *               once by the owner before the first step;
 *   var_multicast : optional NVSwitch multicast mapping covering every rank's
 *               var replica (one store reaches all replicas, in-switch); NULL
 *               makes the pull `world` unicast peer stores;
 *   var_bf16  : optional bf16 [world*shard] replicas, one per rank. When set, the
 *               pull delivers the updated variables rounded to bf16 (nearest-even)
 *               into every rank's var_bf16 — half the NVLink bytes — and the fp32
 *               values stay with the owner only (its shard inside var[rank]).
 *               For workers that compute in bf16 (SURVEY §8 f, rank 2).
 *   stage_multicast : optional NVSwitch multicast mapping covering every rank's
 *               stage buffer; with var_multicast it enables the in-switch form of
 *               the step (ACCEL_OVERLAP_SWITCH below). NULL otherwise.
 *   stage_dtype : ACCEL_F32 (0, the default) — stage slots hold fp32 partial sums.
 *               ACCEL_BF16 — every rank pushes exactly ONE bf16 gradient
 *               (n_workers = 1, grad_dtype = ACCEL_BF16; BASELINE config 4), so the
 *               "partial sum" IS that gradient and travels as raw bf16: half the
 *               reduce-scatter bytes. The owner widens (exact) and sums in rank
 *               order in fp32, so results are bit-identical to fp32 slots. All
 *               ranks of a step must use the same value. */
#define ACCEL_SHARD_FLAG_WORDS 256
typedef struct accel_peers {
  int32_t rank;
  int32_t world;
  uint64_t shard_elems; /* elements per shard (padded arena / world), multiple of 8 */
  float *const *stage;
  float *const *var;
  uint32_t *const *flags;
  float *var_multicast;
  uint16_t *const *var_bf16;
  float *stage_multicast;
  int32_t stage_dtype;
  int32_t reserved0;
} accel_peers;

/* Push (the worker->PS reduce-scatter, fused): sum this rank's n_workers local
 * gradients and store the partial for shard c straight into rank c's stage
 * slot [rank] over NVLink (own shard: local store), then signal every peer.
 * One pass: reads n_workers*sizeof(g), writes 4 B/elem, of which
 * (world-1)/world crosses NVLink. Replaces the gRPC RecvTensor of each gradient
 * to its PS task. [sub_begin, sub_begin+sub_count) selects the
 * same slice of EVERY shard (multiples of 8; sub_count 0 = whole shards), so a
 * step can be cut into slices and pipelined. Every push opens a new epoch,
 * returned in *epoch_out (may be NULL). */
int accel_shard_push(accel_ctx *ctx, const accel_peers *peers, const void *const *grads,
                     int n_workers, int grad_dtype, uint64_t sub_begin, uint64_t sub_count,
                     accel_stream_t stream, uint32_t *epoch_out);

/* Apply + pull (the PS apply and the PS->worker all-gather, fused): wait for all
 * ranks' pushes of `epoch` (0 = the latest push of this rank), sum the `world`
 * stage slots of this rank's shard slice in rank order, apply the optimizer to
 * the slice's state (the var shard lives inside var[rank]), and store the
 * updated values into every rank's var replica over NVLink, then signal.
 * slot0/slot1: this rank's shard-sized optimizer slots. n_contrib = total number
 * of worker gradients summed over all ranks (the divisor for ACCEL_AGG_MEAN).
 * Clipping needs grad_sumsq_dev (the global sum of squares). */
int accel_shard_apply_pull(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                           float *slot1, int n_contrib, const accel_hparams *hp,
                           const float *grad_sumsq_dev, uint64_t sub_begin, uint64_t sub_count,
                           uint32_t epoch, accel_stream_t stream);

/* Block the stream until every rank's push of `epoch` (0 = latest) has landed in
 * this rank's stage slots (device-side, bounded). Needed only when something
 * has to read the slots between push and apply_pull — global-norm clipping:
 * accel_grad_sumsq over the `world` slots of the own shard (they are `world`
 * fp32 "worker" gradients of shard_elems each), an all-reduce of that scalar
 * across ranks, then accel_shard_apply_pull with the result (e). */
int accel_shard_wait_push(accel_ctx *ctx, const accel_peers *peers, uint32_t epoch,
                          accel_stream_t stream);

/* Block the stream until every rank's apply_pull of `epoch` (0 = latest) has
 * landed in this rank's var replica (device-side, bounded wait). */
int accel_shard_wait_pull(accel_ctx *ctx, const accel_peers *peers, uint32_t epoch,
                          accel_stream_t stream);

/* The whole step as one call, pipelined: the arena is cut into n_chunks slices;
 * push of slice k+1 (HBM-bound) runs on `stream` while apply_pull of slice k
 * (NVLink-bound) runs on a context-owned second stream; ends with wait_pull on
 * `stream`. This is the per-step call of a PS replica in a sharded job
 * (what examples/workdir/mnist_replica.py:256 triggers on the PS side). */
int accel_shard_step(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                     float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                     int n_contrib, const accel_hparams *hp, const float *grad_sumsq_dev,
                     int n_chunks, accel_stream_t stream);

/* How the one-kernel step overlaps its phases.
 * INLINE: every CTA pushes slice k, then applies whichever earlier slices have
 *         arrived from all ranks, and blocks only once its own pushes are issued.
 * ROLES : even CTAs only push (per-warp slice tickets, no CTA barrier), odd CTAs
 *         only apply + pull, trailing the pushers by one slice.
 * SWITCH: the in-switch (NVLS) form. Each rank first writes the sum of its local
 *         gradients for the WHOLE arena into its own stage buffer (layout
 *         [world*shard], not slots); the owner of shard r then reads
 *         stage_multicast at its shard with multimem.ld_reduce — the NVSwitch
 *         fetches the `world` local sums and adds them — applies the optimizer,
 *         and writes the result through var_multicast with multimem.st, which
 *         the switch delivers to every replica. Per GPU and direction NVLink
 *         carries (1 + 1/world) of the arena instead of 2(world-1)/world. A
 *         few CTAs reduce+apply+pull while the others sum, slice by slice. The
 *         switch adds in its own order: results match the other forms within
 *         the path's tolerance (rel 1e-5), not bit for bit; with stage_dtype
 *         ACCEL_BF16 the switch returns the sum rounded to bf16 (fp32
 *         accumulation) — within the bf16 path's rel 1e-2. Needs
 *         stage_multicast and var_multicast. When both are NULL the same
 *         protocol runs with `world` peer loads summed in rank order and
 *         `world` peer stores instead (bit-exact; the form a single GPU can run
 *         under accel_shard_step_local_group). Zero-copy: with n_workers = 1 the
 *         caller may pass its own stage buffer as the gradient (grads[0] ==
 *         stage[rank], arena layout, fp32 — or bf16 with stage_dtype ACCEL_BF16):
 *         the worker then writes its gradient where the switch reads it and the
 *         step copies nothing. The other forms reject that aliasing. */
typedef enum accel_overlap {
  ACCEL_OVERLAP_INLINE = 0,
  ACCEL_OVERLAP_ROLES = 1,
  ACCEL_OVERLAP_SWITCH = 2
} accel_overlap;

/* The whole step as ONE cooperative kernel per GPU: push, apply, pull and the
 * closing wait overlapped slice by slice inside the kernel. Same results as
 * accel_shard_step, bit for bit. n_slices in [1,64]. Needs all CTAs resident:
 * the launch is cooperative and fails rather than deadlocks if the GPU cannot
 * hold the grid. */
int accel_shard_step_fused(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                           float *slot1, const void *const *grads, int n_workers, int grad_dtype,
                           int n_contrib, const accel_hparams *hp, const float *grad_sumsq_dev,
                           int n_slices, int overlap, accel_stream_t stream);

/* Development / CI vehicle for the one-kernel step: all n_ranks (<= 8) ranks of a
 * group whose buffers live on ONE GPU are played by one cooperative launch (CTA
 * group g is rank g), so the cross-rank waits are between co-resident CTAs.
 * peers_of_rank[r] is rank r's view; slot0/slot1/grads are per-rank arrays. No
 * reference counterpart. */
int accel_shard_step_local_group(accel_ctx *ctx, const accel_peers *const *peers_of_rank,
                                 int n_ranks, int optimizer, float *const *slot0,
                                 float *const *slot1, const void *const *const *grads,
                                 int n_workers, int grad_dtype, int n_contrib,
                                 const accel_hparams *hp, const float *grad_sumsq_dev,
                                 int n_slices, int overlap, accel_stream_t stream);

/* Synchronise the stream and report whether any cross-GPU wait of this rank
 * timed out (every device-side wait is bounded; a rank that never arrives turns
 * into ACCEL_ERR_STATE here instead of a hung GPU). */
int accel_shard_check(accel_ctx *ctx, const accel_peers *peers, accel_stream_t stream);

/* Recover after accel_shard_check reported a timeout: synchronise the stream,
 * zero this rank's flag words (tickets, epochs, status) and restart its epoch
 * counter at 0. EVERY rank of the group must call it, with a host-side barrier
 * between the last reset and the next step (a peer that still publishes old
 * epochs into freshly zeroed words would be taken for the new step). No
 * reference counterpart (TF-1.4 aborts the session on a lost PS). */
int accel_shard_reset(accel_ctx *ctx, const accel_peers *peers, accel_stream_t stream);

/* Share of the one-kernel step's CTAs that reduce + apply + pull under
 * ACCEL_OVERLAP_SWITCH (percent, 1..90; 0 restores the default: 25 with one
 * local gradient, 8 with several), or that apply + pull under
 * ACCEL_OVERLAP_ROLES (default 50: every second CTA).
 * coop = 1 (default) launches the one-kernel step cooperatively (co-residency
 * guaranteed or the launch fails); coop = 0 uses a plain launch when the grid
 * is at most one CTA per SM — lower launch latency for small shards, safe only
 * while nothing else occupies the GPU (waits stay bounded either way).
 * Benchmark/tuning knob; no reference counterpart. */
int accel_ctx_set_step_shape(accel_ctx *ctx, int reducer_pct, int coop);

/* The sharded step with HOST gradient buffers, per rank (the N > 1 twin of
 * accel_ps_step_host; replaces the gRPC-into-host-memory arrangement of
 * examples/workdir/mnist_replica.py:251-264 on every PS pod of the box). The
 * arena is cut into n_slices slices (the same sub-range of every shard); for
 * slice k the n_workers pinned host gradients are copied to context-owned
 * device buffers over this rank's PCIe link while slice k-1 runs
 * accel_shard_push -> accel_shard_apply_pull on a second stream and slice k-2's
 * part of the updated replica returns to host_var_out (world*shard fp32; may be
 * NULL) on a third. Returns when the step is complete on this rank. All ranks
 * must call it with the same n_slices. Device gradient buffers (n_workers x
 * arena) are allocated by the context at first use and kept. clip_norm must be
 * 0 (the global norm needs the whole gradient first). */
int accel_shard_step_host(accel_ctx *ctx, const accel_peers *peers, int optimizer, float *slot0,
                          float *slot1, const void *const *host_grads, int n_workers,
                          int grad_dtype, int n_contrib, const accel_hparams *hp,
                          float *host_var_out, int n_slices);

#ifdef __cplusplus
} /* extern "C" */
#endif
#endif /* ACCEL_PS_H_ */
