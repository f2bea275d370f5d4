/* b200_backend.h — exported entry points of the ggml backend plugin libggml-b200.so.
 *
 * This is the drop-in boundary: the same symbols and proc-address names a ggml backend exports
 * to ggml's dynamic loader (reference ggml/src/ggml-backend-reg.cpp:194-238 load_backend,
 * ggml/src/ggml-backend-impl.h:215-235).  Unmodified llama.cpp / llama-bench / test-backend-ops
 * load the plugin with   GGML_BACKEND_PATH=/path/to/libggml-b200.so   (ggml-backend-reg.cpp:560-564).
 *
 * The types are ggml's own (opaque here): the plugin is compiled against the reference's
 * ggml/include headers and ggml/src/ggml-backend-impl.h and links the reference's libggml-base.so.
 */
#ifndef B200_BACKEND_H
#define B200_BACKEND_H

#ifdef __cplusplus
extern "C" {
#endif

struct ggml_backend_reg;
struct ggml_backend_buffer_type;
struct ggml_backend_feature;

/* required by the loader: returns the registry {api_version = 2, iface, context}
 * (replaces ggml_backend_cuda_reg, reference ggml-cuda/ggml-cuda.cu:5049-5093 + GGML_BACKEND_DL_IMPL :5117) */
struct ggml_backend_reg * ggml_backend_init(void);

/* optional: 0 = not usable on this system (no sm_100 device), >0 = usable (ggml-backend-impl.h:222-235) */
int ggml_backend_score(void);

/* Names served by reg->iface.get_proc_address (reference ggml-cuda.cu:4943-5046):
 *   "ggml_backend_split_buffer_type"  (int main_device, const float * tensor_split) -> buft
 *        row-split (tensor-parallel) buffer type, replaces ggml_backend_cuda_split_buffer_type (:1065-1107)
 *   "ggml_backend_get_features"       (reg) -> NULL-terminated {name,value} array (:4995-5040)
 *   "ggml_backend_b200_get_stats"     (struct ggml_b200_stats *) -> void      [extension, see below]
 */
struct ggml_backend_buffer_type * ggml_backend_b200_split_buffer_type(int main_device, const float * tensor_split);

/* The row partition of that buffer type as plain arithmetic (no device needed): rows [*row_lo, *row_hi) of `device` for a
 * tensor of `nrows` rows over `n_devices` equal devices (reference get_row_split, ggml-cuda.cu:791-804: cumulative
 * fractions, boundaries rounded down to 128 rows, the last device takes the remainder).  0 on success, -1 on bad arguments. */
int ggml_backend_b200_split_rows(const float * tensor_split, int n_devices, long long nrows, int device, long long * row_lo, long long * row_hi);

/* Extension used by bench.py / tools: counters and device-side timing of graph_compute.
 * Device time is measured with CUDA events recorded on the backend stream around every graph
 * launch when the environment variable GGML_B200_TIMING=1 is set. */
struct ggml_b200_stats {
    unsigned long long graphs_computed;   /* graph_compute calls */
    unsigned long long graphs_replayed;   /* of which served by a CUDA-graph replay */
    unsigned long long kernels_launched;  /* kernel launches, counting every node of every replay */
    unsigned long long nodes_last_graph;  /* kernel nodes in the most recent graph */
    double device_ms_total;               /* sum of event-timed graph durations (GGML_B200_TIMING=1) */
    double device_ms_last;
    double host_ms_total;                 /* wall time spent inside graph_compute (host side: keying, planning, launches) */
};
void ggml_backend_b200_get_stats(struct ggml_b200_stats * out);
void ggml_backend_b200_reset_stats(void);

#ifdef __cplusplus
}
#endif
#endif
