/* kungfu_b200 host runtime — C ABI (libkungfu_host.so).
 *
 * One kf_peer per process. Functions return 0 on success unless stated; the
 * text of the last error is kf_last_error(). dtype codes are KungFu_Datatype
 * compatible (srcs/cpp/include/kungfu/dtype.h:21-39) plus bf16 = (4<<16|2<<8|8);
 * op: 0 sum, 1 min, 2 max, 3 prod (srcs/cpp/include/kungfu/op.h:8-13).
 * Reference surface: /root/reference/srcs/go/libkungfu-comm/{main.go,
 * collective.go,adapt.go,monitoring.go}, /root/reference/srcs/cpp/include/
 * kungfu/python/init.h:4-31.
 */
#ifndef KUNGFU_HOST_H
#define KUNGFU_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct kf_peer kf_peer;

const char *kf_last_error(void);
kf_peer *kf_peer_create_from_env(void);
int kf_peer_start(kf_peer *p);
void kf_peer_close(kf_peer *p);

int kf_rank(kf_peer *p);
int kf_size(kf_peer *p);
int kf_local_rank(kf_peer *p);
int kf_local_size(kf_peer *p);
int kf_host_count(kf_peer *p);
int kf_detached(kf_peer *p);
int kf_cluster_version(kf_peer *p);
uint64_t kf_uid(kf_peer *p);
int kf_peer_addr(kf_peer *p, int rank, char *buf, int cap);

int kf_barrier(kf_peer *p);
int kf_all_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name);
int kf_monitored_all_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op,
                            const int32_t *tree, int tree_len, const char *name);
/* one all-reduce along the given tree (father array; empty = active strategy) without installing it
 * (Session.AllReduceWith, /root/reference/srcs/go/kungfu/session/allreduce.go:15-27) */
int kf_all_reduce_with(kf_peer *p, const int32_t *tree, int tree_len, const void *send, void *recv, int64_t count,
                       int dtype, int op, const char *name);
int kf_cross_all_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name);
int kf_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name);
int kf_broadcast(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name);
int kf_local_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name);
int kf_local_broadcast(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name);
int kf_all_gather(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name);
int kf_gather(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name);
int kf_consensus(kf_peer *p, const void *data, int64_t nbytes, const char *name, int *agree);

/* async: returns a handle (>0) or -1; kf_wait_handle returns 0 iff the op succeeded */
int64_t kf_all_reduce_async(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name);
int64_t kf_broadcast_async(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name);
int kf_wait_handle(kf_peer *p, int64_t handle);
/* Run fn(arg) on the runtime's worker threads, then on_done(arg, rc) on the same thread (may be NULL).
 * This is the path behind the callback overloads of kungfu::Peer (the reference ran them on the Go
 * runtime's goroutines, srcs/go/libkungfu-comm/main.go:163-187). kf_peer_close() runs every job
 * that was queued before it and joins the workers, so no job outlives the peer. */
int kf_submit(kf_peer *p, int (*fn)(void *), void *arg, void (*on_done)(void *, int));

/* p2p model store; version may be NULL (unversioned). kf_request returns 0 iff found. */
int kf_save(kf_peer *p, const char *version, const char *name, const void *data, int64_t nbytes);
int kf_request(kf_peer *p, int target_rank, const char *version, const char *name, void *buf, int64_t nbytes);

/* async request: a handle for kf_wait_handle (0 iff the blob was found with that size) */
int64_t kf_request_async(kf_peer *p, int target_rank, const char *version, const char *name, void *buf, int64_t nbytes);

/* Double-buffered pull of a peer's fused model (the reference's ModelBuffer + AsyncModelAveraging /
 * AsyncRequestModel, srcs/cpp/src/tensorflow/ops/cpu/peer_to_peer.cpp:156-258,411-524): the first
 * kf_prefetch_next blocks on the request; every call then starts the request for the NEXT call on the
 * runtime's workers (unless one is still in flight) and hands out the newest complete model, so the
 * training loop never waits on the network. *fresh (may be NULL) = number of completed prefetches so far.
 * kf_prefetch_free waits for the request in flight. */
typedef struct kf_prefetch kf_prefetch;
kf_prefetch *kf_prefetch_new(kf_peer *p, const char *name, int64_t nbytes);
int kf_prefetch_next(kf_prefetch *pf, int target_rank, void *out, int64_t *fresh);
void kf_prefetch_free(kf_prefetch *pf);

/* Fused-model helpers for a model that lives in n separate tensors (fp32 unless noted), so the host path touches
 * every byte once instead of concatenating, copying and splitting in the caller:
 *   kf_save_parts        gather n byte ranges straight into a new blob of the store (any dtype);
 *   kf_blend_f32         t_i <- 0.5 * (t_i + other[off_i ...]) for every tensor, 8 lanes at a time;
 *   kf_prefetch_blend_f32  kf_prefetch_next + kf_blend_f32 without the intermediate copy: the newest complete
 *                          model is blended into the tensors directly from the double buffer.
 * counts are elements for the f32 calls, bytes for kf_save_parts. */
int kf_save_parts(kf_peer *p, const char *version, const char *name, const void *const *parts, const int64_t *nbytes, int n);
void kf_blend_f32(float *const *tensors, const int64_t *counts, int n, const float *other);
int kf_prefetch_blend_f32(kf_prefetch *pf, int target_rank, float *const *tensors, const int64_t *counts, int n, int64_t *fresh);

/* elastic control */
int kf_resize_cluster(kf_peer *p, int new_size, int *changed, int *detached);
int kf_resize_cluster_from_url(kf_peer *p, int *changed, int *detached);
/* *differs = 1 if the config server publishes a cluster other than the current one (same fetch +
 * consensus as above, nothing is changed). Collective. */
int kf_peek_cluster_change(kf_peer *p, int *differs);
int kf_propose_new_size(kf_peer *p, int new_size);

/* adaptation / monitoring */
int kf_set_tree(kf_peer *p, const int32_t *forest, int n);
int kf_set_strategy(kf_peer *p, const char *name);
/* active global strategies as text, one "reduce=[n]{(i)(i->j)...} bcast=[n]{...}" line per graph pair */
int kf_current_strategies(kf_peer *p, char *out, int cap);
int kf_current_strategy_name(kf_peer *p, char *out, int cap);
int kf_calc_stats(kf_peer *p);
int kf_log_stats(kf_peer *p);
int kf_print_strategy_stats(kf_peer *p);
int kf_check_interference(kf_peer *p); /* 1 = majority voted to switch */
int kf_get_peer_latencies(kf_peer *p, float *out, int n);
int kf_egress_rates(kf_peer *p, float *out, int n);

/* descriptor passing between colocated workers (CUDA VMM handles) */
int kf_send_fd(kf_peer *p, int target_rank, const char *tag, int fd);
int kf_recv_fd(kf_peer *p, int src_rank, const char *tag, int *fd);

/* planner utilities exposed for tests / tools */
int kf_gen_strategy_graphs(const char *peers_csv, const char *strategy, char *out, int cap);
/* This rank's view of the active strategies, compiled for the device executor (kfb_graph_allreduce):
 * out = {n_plans, n_rounds, n_reduce_rounds, n_edges, round_begin[n_rounds + 1], edge_plan[n_edges],
 * edge_src[n_edges]}; returns the number of ints written or -1. kf_device_plan_for compiles a named
 * strategy over an explicit peer list (tests, tools). */
int kf_device_plan(kf_peer *p, int32_t *out, int cap);
int kf_device_plan_for(const char *peers_csv, const char *strategy, int rank, int32_t *out, int cap);
int kf_minimum_spanning_tree(const float *weights, int n, int32_t *edges_out);
void kf_transform2(const void *x, const void *y, void *z, int64_t n, int dtype, int op);
int kf_cluster_resize(const char *cluster_json, int new_size, char *out, int cap);
int kf_parse_hostfile(const char *text, char *out, int cap);
int kf_gen_peer_list(const char *hostlist, int np, const char *port_range, char *out, int cap);

/* tools */
int kf_run_main(int argc, char **argv);            /* kungfu-run */
int kf_config_server_main(int argc, char **argv);  /* kungfu-config-server */

#ifdef __cplusplus
}
#endif
#endif
