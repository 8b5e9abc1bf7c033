/* C ABI of libttd_io.so — host-side I/O of tensorflow_train_distributed_b200 (no CUDA, no TensorFlow).
 *
 * Functions return 0 on success (or a non-NULL handle); on failure a non-zero code / NULL and
 * ttdio_last_error() explains. Citations: /root/reference/distribute_training.py and SURVEY.md A.4.
 */
#ifndef TTD_IO_H_
#define TTD_IO_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* ttdio_last_error(void);

/* CRC-32C (Castagnoli); *_masked applies TensorFlow's rotate-and-add mask. */
uint32_t ttdio_crc32c(const void* data, uint64_t n);
uint32_t ttdio_crc32c_software(const void* data, uint64_t n); /* table path, for cross-checking the SSE4.2 path */
uint32_t ttdio_crc32c_masked(const void* data, uint64_t n);

/* TF DataType codes used in BundleEntryProto.dtype */
enum { TTDIO_DT_FLOAT = 1, TTDIO_DT_DOUBLE = 2, TTDIO_DT_INT32 = 3, TTDIO_DT_UINT8 = 4, TTDIO_DT_INT64 = 9,
       TTDIO_DT_BOOL = 10, TTDIO_DT_BFLOAT16 = 14, TTDIO_DT_HALF = 19 };

/* ---- tensor bundle ("V2 checkpoint": what Saver writes for checkpoint_dir, reference :209-213) ---- */
/* Phase 1 (every shard owner): write <prefix>.data-<shard>-of-<num> and a partial index under <prefix>_temp/. */
int ttdio_bundle_write_shard(const char* prefix, int shard_id, int num_shards, int n_tensors,
                             const char* const* names, const int32_t* dtypes, const int32_t* ndims,
                             const int64_t* const* shapes, const void* const* data, const int64_t* nbytes);
/* Phase 2 (chief): merge the partial indices into <prefix>.index (MergeV2Checkpoints). */
int ttdio_bundle_merge(const char* prefix, int num_shards);

void* ttdio_bundle_open(const char* prefix); /* parses and checksums <prefix>.index */
void ttdio_bundle_close(void* reader);
int ttdio_bundle_num_entries(void* reader);
int ttdio_bundle_num_shards(void* reader);
/* i-th entry in key order; shape8 must hold 8 values; name stays valid until close. */
int ttdio_bundle_entry(void* reader, int i, const char** name, int32_t* dtype, int32_t* ndim, int64_t* shape8,
                       int32_t* shard_id, int64_t* offset, int64_t* size, uint32_t* crc_masked);
/* Reads one tensor's bytes (nbytes must equal its stored size) and verifies its checksum. */
int ttdio_bundle_read(void* reader, const char* name, void* out, int64_t nbytes);

/* ---- TensorBoard events (tf.summary.scalar, reference :128-132) ------------------------------- */
void* ttdio_events_open(const char* logdir, double wall_time, const char* filename_suffix);
const char* ttdio_events_path(void* writer);
int ttdio_events_add_scalar(void* writer, const char* tag, float value, int64_t step, double wall_time);
int ttdio_events_flush(void* writer);
void ttdio_events_close(void* writer);

/* ---- multi-thread batch reader (README.md:5-7 "ReadByQueue"; single-thread form: next_batch, :224) -------- */
/* Worker threads gather rows of images[n][feat] / labels[n] into a ring of caller-provided batch buffers
 * (x_slots[i]: batch*feat floats, y_slots[i]: batch int64; pinned host memory on a GPU box). The example order
 * comes from the caller: batch k = positions [k*batch, (k+1)*batch) of the concatenation of the pushed epoch
 * permutations (= next_batch's roll-over). Batches come back in order. */
void* ttdio_loader_create(const float* images, const int64_t* labels, int64_t n, int64_t feat, int32_t batch,
                          int32_t ring, int32_t threads, float* const* x_slots, int64_t* const* y_slots);
int ttdio_loader_push_epoch(void* loader, const int64_t* perm, int64_t count);
int64_t ttdio_loader_backlog(void* loader);              /* pushed stream positions not yet consumed */
int ttdio_loader_next(void* loader, int32_t timeout_ms); /* ring slot of the next batch; -1 timeout; -2 closed */
int ttdio_loader_release(void* loader, int64_t upto);    /* batches 0..upto may be overwritten */
void ttdio_loader_destroy(void* loader);

/* ---- parameter-server task (--job_name=ps, reference :179-181; in the reference this is TensorFlow's C++ gRPC
 * worker service). One listening TCP endpoint per ps; request = {op, a, b, c} int64 + optional float32 payload.
 * Ops: 0 READY, 1 INIT, 2 PULL, 3 PUSH_ASYNC, 4 PUSH_SYNC, 5 TOKEN, 6 GET_STATE, 7 DONE, 8 SYNC_INIT, 9 HELLO
 * (see csrc/io/ps_server.cc for the payloads and the SyncReplicasOptimizer semantics). */
typedef struct ttdps_config {
  int64_t numel;               /* floats in this ps's shard (variables placed round-robin, reference :186-188) */
  int32_t n_workers, n_agg;    /* workers in the cluster; replicas_to_aggregate (:146) */
  int32_t tokens_per_step;     /* tokens enqueued per global step: max(total_num_replicas, n_agg) (:147) */
  int32_t owns_global_step;    /* 1 on the ps that holds global_step (ps 0) */
  int32_t reg_kind;            /* 0 none, 1 L1, 2 L2 */
  int32_t use_ema;
  float reg_scale, ema_decay;
  double idle_timeout_secs;    /* > 0: serve() returns 1 after this long without a request (once initialised) */
} ttdps_config;
void* ttdps_server_create(const char* bind_host, int32_t port, const void* config, const float* reg_mask);
int32_t ttdps_server_port(void* server);
int ttdps_server_serve(void* server); /* blocks until every worker sent DONE (0) or the idle limit passed (1) */
void ttdps_server_stats(void* server, int64_t* out6); /* applied, dropped_stale, pulls, global_step, initialized, done */
void ttdps_server_destroy(void* server);

#ifdef __cplusplus
}
#endif
#endif /* TTD_IO_H_ */
