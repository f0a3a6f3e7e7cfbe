// Input loading: the text grammar of the reference driver, parsed straight into
// flat (optionally pinned) host buffers.
//
// Grammar (reference: common.cpp:93-117, generate_input.py:11-21):
//   line 1          : N Q A
//   next N lines    : label a_0 ... a_{A-1}
//   next Q lines    : Q k a_0 ... a_{A-1}      (the leading 'Q' is mandatory)
// The reference parses with one stringstream per line on rank 0, outside its
// timer.  Here the whole stream is slurped once, newline positions are indexed,
// and lines are converted by a pool of threads with std::from_chars (correctly
// rounded, so the doubles are bit-identical to the reference's extraction).
#ifndef DMLP_IO_H_
#define DMLP_IO_H_

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "dmlp/views.h"

namespace dmlp {

/// Allocation hooks so a GPU backend can ask for page-locked memory
/// (cudaHostAlloc) while CPU-only builds use the C heap.
struct HostAllocator {
  void *(*alloc)(size_t bytes) = nullptr;
  void (*release)(void *p) = nullptr;
};
HostAllocator malloc_allocator();

/// Owner of the flat problem arrays.  Movable, not copyable.
class Problem {
 public:
  Problem() = default;
  explicit Problem(HostAllocator a) : alloc_(a) {}
  ~Problem();
  Problem(const Problem &) = delete;
  Problem &operator=(const Problem &) = delete;
  Problem(Problem &&o) noexcept;
  Problem &operator=(Problem &&o) noexcept;

  /// (Re)allocate for the given shape.
  void resize(int64_t n, int64_t q, int a);

  int64_t num_data = 0, num_queries = 0;
  int num_attrs = 0;
  double *data_attrs = nullptr;   ///< n x a
  int *data_labels = nullptr;     ///< n
  double *query_attrs = nullptr;  ///< q x a
  int *query_k = nullptr;         ///< q
  int32_t *data_attrs_i32 = nullptr;   ///< n x a, set by encode_decimal6
  int32_t *query_attrs_i32 = nullptr;  ///< q x a

  /// The loader's second output: every attribute as the integer i with value == double(i) / 1e6
  /// (what the text grammar's "%.6f" holds).  Returns false, and keeps no integer form, when some
  /// value is not such a quotient or does not fit 32 bits.  `threads` <= 0: hardware concurrency.
  bool encode_decimal6(int threads);

  DatasetView dataset() const {
    DatasetView v; v.n = num_data; v.a = num_attrs; v.attrs = data_attrs;
    v.labels = data_labels; v.id_base = 0; v.attrs_i32 = data_attrs_i32; return v;
  }
  QueryView queries() const {
    QueryView v; v.q = num_queries; v.a = num_attrs; v.attrs = query_attrs;
    v.k = query_k; v.id_base = 0; v.attrs_i32 = query_attrs_i32; return v;
  }

 private:
  void free_all();
  HostAllocator alloc_ = malloc_allocator();
};

/// Parse a complete input held in memory.  Throws std::runtime_error with the
/// reference's messages ("Line is empty", "Line is wrongly formatted") on the
/// same conditions the reference checks.  `threads` <= 0 picks the hardware
/// concurrency.
void parse_problem(const char *text, size_t len, int threads, Problem *out);

/// One update record `id v0 v1 ...` (reference: parse_update, common.cpp:46-55 -- never
/// called there either; kept so the Update type of common.h has its parser).
bool parse_update_line(const char *text, size_t len, int *id, std::vector<double> *new_attrs);

/// Read everything from `f` (stdin in the driver) and parse it.
void read_problem(std::FILE *f, int threads, Problem *out);

/// Deterministic synthetic problem with the generator's distribution
/// (reference: generate_input.py:13-21): labels uniform in [0, num_labels),
/// attributes uniform in [lo, hi] rounded to six decimals, k uniform in
/// [min_k, min(max_k, n)].  This is the host twin of the device fill kernel and
/// uses the same counter-based hash, so both produce identical problems.
struct SynthSpec {
  int64_t n = 0, q = 0;
  int a = 0;
  double lo = -100.0, hi = 100.0;
  int min_k = 1, max_k = 1, num_labels = 2;
  uint64_t seed = 42;
};
void synth_problem(const SynthSpec &spec, Problem *out);

/// Counter-based generator shared by host and device fill code.
/// Stream ids: 0 = data attrs, 1 = labels, 2 = query attrs, 3 = query k.
inline uint64_t synth_hash(uint64_t seed, uint64_t stream, uint64_t index) {
  uint64_t z = seed * 0x9E3779B97F4A7C15ULL + stream * 0xD1B54A32D192ED03ULL + index;
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
/// Attribute value for hash `h`: uniform in [lo, hi], six decimals, as the
/// double a text round trip through "%.6f" would give.
inline double synth_attr(uint64_t h, double lo, double hi) {
  double u = static_cast<double>(h >> 11) * (1.0 / 9007199254740992.0);  // [0,1)
  double x = lo + (hi - lo) * u;
  double scaled = x * 1e6;
  long long i = static_cast<long long>(scaled < 0 ? scaled - 0.5 : scaled + 0.5);
  return static_cast<double>(i) / 1e6;
}

/// values -> integers i with value == double(i) / 1e6 exactly (the six decimals of the text grammar);
/// false when some value is not such a quotient or does not fit 32 bits (`out` is then undefined).
bool encode_decimal6_values(const double *values, int64_t count, int threads, int32_t *out);

/// Bit-packed decimal6 wire format (the device side is dmlp_pack_bits / dmlp_unpack_bits,
/// src/kernels/pack.cu): value i travels as the `bits`-bit number i + 2^(bits-1), a row of `a`
/// values is a little-endian bit stream padded to `words` 32-bit words.  decimal6_bits: the
/// smallest width that holds every value (values in [-100, 100] need 28).
int decimal6_bits(const int32_t *values, int64_t count);
void pack_bits_rows(const int32_t *src, int64_t rows, int a, int bits, int words, uint32_t *dst);
void unpack_bits_rows(const uint32_t *src, int64_t rows, int a, int bits, int words, int32_t *dst);

/// Smallest and largest attribute value over data and queries (threads <= 0: hardware
/// concurrency); 0, 0 for an empty problem.
void value_range(const Problem &p, int threads, double *lo, double *hi);

/// Write a Problem in the reference's text grammar (six decimals).
void write_problem_text(const Problem &p, std::FILE *f);

}  // namespace dmlp

#endif  // DMLP_IO_H_
