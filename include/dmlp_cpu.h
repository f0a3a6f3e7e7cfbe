/* C ABI of libdmlp_cpu.so: the CPU oracle / multi-"rank" strategies, the text
 * parser and the host synthetic generator, for ctypes and other FFI users.
 * All functions return 0 on success and a negative value on failure;
 * dmlp_cpu_last_error() then describes the failure (thread-local). */
#ifndef DMLP_CPU_C_H_
#define DMLP_CPU_C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char *dmlp_cpu_last_error(void);

/* Whole job on the CPU (reference: Engine::KNN, engine.cpp:20-350).
 * strategy: "serial" | "grid" | "gather" | "tree" | "farm" | "kdtree".
 * out_ids / out_dist hold sum(max(k_i,0)) entries in query order (CSR by k). */
int dmlp_cpu_knn(const double *data_attrs, const int *data_labels, int64_t n,
                 const double *query_attrs, const int *query_k, int64_t q, int a,
                 const char *strategy, int ranks, int rows, int cols, int threads,
                 int *out_label, uint64_t *out_checksum, int *out_ids, double *out_dist);

/* Exact squared distance with the reference rounding (engine.cpp:12-18). */
double dmlp_exact_sqdist(const double *a, const double *b, int dim);

/* reportResult's fold (common.cpp:59-68). */
uint64_t dmlp_result_checksum(int label, const int *ids, int64_t k);

/* Merge operator of the tree strategy: two sorted k-lists -> first k. */
void dmlp_merge_topk(const double *xd, const int *xi, const double *yd, const int *yi, int k,
                     double *od, int *oi);

/* The two halves of the distributed job, for multi-process CPU runs (gloo): per-shard
 * sorted, sentinel-padded k-lists with global ids (engine.cpp:233-257), and the
 * owner-side merge + vote + checksum over `sources` such lists (engine.cpp:289-347).
 * Lists are [q][kmax] with stride kmax. */
int dmlp_cpu_local_lists(const double *data_attrs, const int *data_labels, int64_t n, int id_base,
                         const double *query_attrs, const int *query_k, int64_t q, int a, int kmax, int threads,
                         double *out_dist, int *out_id, int *out_label);
int dmlp_cpu_finalize_lists(const double *dist, const int *id, const int *label, int sources, int64_t q, int kmax,
                            const int *query_k, int *out_label, uint64_t *out_checksum);

/* Text parser (common.cpp:93-117 grammar).  Two steps: read the header, then
 * fill caller-owned arrays. */
int dmlp_parse_header(const char *text, size_t len, int64_t *n, int64_t *q, int *a);
int dmlp_parse_text(const char *text, size_t len, int threads, double *data_attrs, int *data_labels,
                    double *query_attrs, int *query_k);

/* The loader's integer form of six-decimal inputs: out[i] = values[i] * 1e6 as a 32-bit integer,
 * *ok = 1 when every value equals double(out[i]) / 1e6 exactly (then a GPU backend can move 4
 * bytes per value over PCIe and rebuild the doubles bit for bit), 0 otherwise. */
int dmlp_encode_decimal6_host(const double *values, int64_t count, int threads, int32_t *out, int *ok);
/* The bit-packed form of those integers, as a host loader produces it (the device side is
 * dmlp_pack_bits / dmlp_unpack_bits in dmlp_knn.h): `bits` per value (dmlp_decimal6_bits_host: the
 * smallest width for a set of values), rows padded to `words` >= ceil(a * bits / 32) 32-bit words. */
int dmlp_decimal6_bits_host(const int32_t *values, int64_t count, int *bits);
int dmlp_pack_bits_host(const int32_t *src, int64_t rows, int a, int bits, int words, uint32_t *dst);
int dmlp_unpack_bits_host(const uint32_t *src, int64_t rows, int a, int bits, int words, int32_t *dst);

/* One update record "id v0 v1 ..." (reference: parse_update, common.cpp:46-55; unused by its
 * engine).  *count receives the number of values found; at most `capacity` are stored. */
int dmlp_parse_update(const char *text, size_t len, int *id, double *attrs, int capacity, int *count);

/* Host twin of the device fill kernel (generate_input.py:13-21 distribution). */
int dmlp_synth_fill(int64_t n, int64_t q, int a, double lo, double hi, int min_k, int max_k,
                    int num_labels, uint64_t seed, double *data_attrs, int *data_labels,
                    double *query_attrs, int *query_k);

/* Grid helpers shared with the Python parallel layer. */
void dmlp_choose_grid(int ranks, int *rows, int *cols);
int64_t dmlp_block_begin(int64_t total, int parts, int p);
int64_t dmlp_block_count(int64_t total, int parts, int p);

#ifdef __cplusplus
}
#endif
#endif /* DMLP_CPU_C_H_ */
