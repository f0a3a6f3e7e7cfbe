// Text loader, synthetic generator and text writer.  See include/dmlp/io.h.
#include "dmlp/io.h"

#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace dmlp {

HostAllocator malloc_allocator() {
  HostAllocator a;
  a.alloc = [](size_t bytes) -> void * { return std::malloc(bytes ? bytes : 1); };
  a.release = [](void *p) { std::free(p); };
  return a;
}

Problem::~Problem() { free_all(); }

void Problem::free_all() {
  if (alloc_.release) {
    if (data_attrs) alloc_.release(data_attrs);
    if (data_labels) alloc_.release(data_labels);
    if (query_attrs) alloc_.release(query_attrs);
    if (query_k) alloc_.release(query_k);
    if (data_attrs_i32) alloc_.release(data_attrs_i32);
    if (query_attrs_i32) alloc_.release(query_attrs_i32);
  }
  data_attrs = query_attrs = nullptr;
  data_labels = query_k = nullptr;
  data_attrs_i32 = query_attrs_i32 = nullptr;
}

Problem::Problem(Problem &&o) noexcept { *this = std::move(o); }

Problem &Problem::operator=(Problem &&o) noexcept {
  if (this != &o) {
    free_all();
    alloc_ = o.alloc_;
    num_data = o.num_data; num_queries = o.num_queries; num_attrs = o.num_attrs;
    data_attrs = o.data_attrs; data_labels = o.data_labels;
    query_attrs = o.query_attrs; query_k = o.query_k;
    data_attrs_i32 = o.data_attrs_i32; query_attrs_i32 = o.query_attrs_i32;
    o.data_attrs = o.query_attrs = nullptr;
    o.data_labels = o.query_k = nullptr;
    o.data_attrs_i32 = o.query_attrs_i32 = nullptr;
    o.num_data = o.num_queries = 0; o.num_attrs = 0;
  }
  return *this;
}

void Problem::resize(int64_t n, int64_t q, int a) {
  free_all();
  if (n < 0 || q < 0 || a < 0) throw std::runtime_error("negative problem dimension");
  num_data = n; num_queries = q; num_attrs = a;
  auto get = [&](size_t bytes) {
    void *p = alloc_.alloc(bytes);
    if (!p) throw std::bad_alloc();
    return p;
  };
  data_attrs = static_cast<double *>(get(sizeof(double) * size_t(n) * size_t(a)));
  data_labels = static_cast<int *>(get(sizeof(int) * size_t(n)));
  query_attrs = static_cast<double *>(get(sizeof(double) * size_t(q) * size_t(a)));
  query_k = static_cast<int *>(get(sizeof(int) * size_t(q)));
}

namespace {

inline const char *skip_blank(const char *p, const char *end) {
  while (p < end && (*p == ' ' || *p == '\t' || *p == '\r')) ++p;
  return p;
}

// One whitespace-delimited number; a missing or malformed token yields 0, the
// value a failed stream extraction leaves behind in the reference's parser.
inline const char *take_double(const char *p, const char *end, double *v) {
  p = skip_blank(p, end);
  if (p < end && *p == '+') ++p;
  auto r = std::from_chars(p, end, *v);
  if (r.ec != std::errc()) { *v = 0.0; return end; }
  return r.ptr;
}

inline const char *take_int(const char *p, const char *end, int *v) {
  p = skip_blank(p, end);
  if (p < end && *p == '+') ++p;
  auto r = std::from_chars(p, end, *v);
  if (r.ec != std::errc()) { *v = 0; return end; }
  return r.ptr;
}

struct Line { const char *b; const char *e; };

}  // namespace

void parse_problem(const char *text, size_t len, int threads, Problem *out) {
  const char *end = text + len;
  // header
  const char *nl = static_cast<const char *>(std::memchr(text, '\n', len));
  const char *hend = nl ? nl : end;
  int n = 0, q = 0, a = 0;
  const char *p = take_int(text, hend, &n);
  p = take_int(p, hend, &q);
  take_int(p, hend, &a);
  out->resize(n, q, a);

  // index the N + Q payload lines
  const size_t want = size_t(n) + size_t(q);
  std::vector<Line> lines;
  lines.reserve(want);
  const char *cur = nl ? nl + 1 : end;
  while (lines.size() < want) {
    if (cur >= end) { lines.push_back(Line{end, end}); continue; }
    const char *e = static_cast<const char *>(std::memchr(cur, '\n', size_t(end - cur)));
    if (!e) e = end;
    lines.push_back(Line{cur, e});
    cur = e + 1;
  }

  if (threads <= 0) threads = int(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  if (size_t(threads) > want) threads = int(want ? want : 1);

  std::vector<std::string> errors;
  errors.resize(static_cast<size_t>(threads));
  auto work = [&](int t) {
    size_t b = want * size_t(t) / size_t(threads), e = want * size_t(t + 1) / size_t(threads);
    for (size_t i = b; i < e; ++i) {
      const char *s = lines[i].b, *le = lines[i].e;
      while (le > s && (le[-1] == '\r')) --le;
      if (i < size_t(n)) {
        if (s == le) { errors[t] = "Line is empty"; return; }
        s = take_int(s, le, &out->data_labels[i]);
        double *row = out->data_attrs + i * size_t(a);
        for (int j = 0; j < a; ++j) s = take_double(s, le, &row[j]);
      } else {
        size_t qi = i - size_t(n);
        if (s == le || *s != 'Q') {
          errors[t] = "Line is wrongly formatted";
          return;
        }
        ++s;
        s = take_int(s, le, &out->query_k[qi]);
        double *row = out->query_attrs + qi * size_t(a);
        for (int j = 0; j < a; ++j) s = take_double(s, le, &row[j]);
      }
    }
  };
  if (threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
    for (auto &th : pool) th.join();
  }
  for (auto &e : errors)
    if (!e.empty()) throw std::runtime_error(e);
}

bool parse_update_line(const char *text, size_t len, int *id, std::vector<double> *new_attrs) {
  const char *end = text + len;
  while (end > text && (end[-1] == '\n' || end[-1] == '\r')) --end;
  const char *p = skip_blank(text, end);
  if (p == end) return false;
  p = take_int(p, end, id);
  new_attrs->clear();
  for (;;) {
    p = skip_blank(p, end);
    if (p < end && *p == '+') ++p;
    if (p >= end) break;
    double v = 0.0;
    auto r = std::from_chars(p, end, v);
    if (r.ec != std::errc()) break;  // a failed extraction ends the list, as operator>> does
    new_attrs->push_back(v);
    p = r.ptr;
  }
  return true;
}

void read_problem(std::FILE *f, int threads, Problem *out) {
  std::vector<char> buf;
  size_t used = 0;
  buf.resize(1 << 22);
  for (;;) {
    if (used == buf.size()) buf.resize(buf.size() * 2);
    size_t got = std::fread(buf.data() + used, 1, buf.size() - used, f);
    if (got == 0) break;
    used += got;
  }
  parse_problem(buf.data(), used, threads, out);
}

bool encode_decimal6_values(const double *v, int64_t n, int threads, int32_t *out) {
  if (threads <= 0) threads = int(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  if (n < (1 << 16)) threads = 1;
  std::vector<char> bad(size_t(threads), 0);
  auto work = [&](int t) {
    for (int64_t i = n * t / threads, e = n * (t + 1) / threads; i < e; ++i) {
      const double scaled = v[i] * 1e6;
      if (!(scaled >= -2147483647.0 && scaled <= 2147483647.0)) { bad[size_t(t)] = 1; return; }   // (also rejects NaN)
      const long long k = std::llrint(scaled);
      // the device rebuilds the correctly rounded quotient k / 1e6: it must be this very double
      // (-0.0 comes back as +0.0: no difference or square can tell them apart)
      if (static_cast<double>(k) / 1e6 != v[i]) { bad[size_t(t)] = 1; return; }
      out[i] = static_cast<int32_t>(k);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto &th : pool) th.join();
  for (char b : bad)
    if (b) return false;
  return true;
}

int decimal6_bits(const int32_t *v, int64_t n) {
  long long big = 0;
  for (int64_t i = 0; i < n; ++i) {
    const long long m = v[i] < 0 ? -static_cast<long long>(v[i]) : v[i];
    big = m > big ? m : big;
  }
  int bits = 2;   // i + 2^(bits-1) must fit `bits` bits: bits = bit_length(max |i|) + 1
  while (bits < 32 && (1LL << (bits - 1)) <= big) ++bits;
  return bits;
}

void pack_bits_rows(const int32_t *src, int64_t rows, int a, int bits, int words, uint32_t *dst) {
  if (bits < 2 || bits > 32 || static_cast<long long>(words) * 32 < static_cast<long long>(a) * bits)
    throw std::invalid_argument("pack_bits: bits must be in [2, 32] and the row must fit its words");
  const uint64_t mask = bits == 32 ? 0xFFFFFFFFULL : ((1ULL << bits) - 1ULL);
  const uint32_t bias = 1u << (bits - 1);
  for (int64_t r = 0; r < rows; ++r) {
    const int32_t *row = src + r * a;
    uint32_t *out = dst + r * words;
    for (int w = 0; w < words; ++w) out[w] = 0;
    uint64_t acc = 0;      // bits not yet written, little end first
    int have = 0, w = 0;
    for (int c = 0; c < a; ++c) {
      acc |= ((static_cast<uint64_t>(static_cast<uint32_t>(row[c]) + bias)) & mask) << have;
      have += bits;
      while (have >= 32) {
        out[w++] = static_cast<uint32_t>(acc);
        acc >>= 32;
        have -= 32;
      }
    }
    if (have > 0) out[w] = static_cast<uint32_t>(acc);
  }
}

void unpack_bits_rows(const uint32_t *src, int64_t rows, int a, int bits, int words, int32_t *dst) {
  if (bits < 2 || bits > 32 || static_cast<long long>(words) * 32 < static_cast<long long>(a) * bits)
    throw std::invalid_argument("unpack_bits: bits must be in [2, 32] and the row must fit its words");
  const uint64_t mask = bits == 32 ? 0xFFFFFFFFULL : ((1ULL << bits) - 1ULL);
  const uint32_t bias = 1u << (bits - 1);
  for (int64_t r = 0; r < rows; ++r) {
    const uint32_t *row = src + r * words;
    for (int c = 0; c < a; ++c) {
      const long long off = static_cast<long long>(c) * bits;
      const int w = static_cast<int>(off >> 5), sh = static_cast<int>(off & 31);
      uint64_t v = row[w];
      if (sh + bits > 32) v |= static_cast<uint64_t>(row[w + 1]) << 32;
      dst[r * a + c] = static_cast<int32_t>(static_cast<uint32_t>((v >> sh) & mask) - bias);
    }
  }
}

bool Problem::encode_decimal6(int threads) {
  if (alloc_.release) {
    if (data_attrs_i32) alloc_.release(data_attrs_i32);
    if (query_attrs_i32) alloc_.release(query_attrs_i32);
  }
  data_attrs_i32 = query_attrs_i32 = nullptr;
  const int64_t nd = num_data * num_attrs, nq = num_queries * num_attrs;
  if (nd + nq == 0) return false;
  int32_t *xi = static_cast<int32_t *>(alloc_.alloc(sizeof(int32_t) * size_t(nd > 0 ? nd : 1)));
  int32_t *qi = static_cast<int32_t *>(alloc_.alloc(sizeof(int32_t) * size_t(nq > 0 ? nq : 1)));
  const bool ok = xi != nullptr && qi != nullptr && encode_decimal6_values(data_attrs, nd, threads, xi) &&
                  encode_decimal6_values(query_attrs, nq, threads, qi);
  if (!ok) {
    if (alloc_.release) {
      if (xi) alloc_.release(xi);
      if (qi) alloc_.release(qi);
    }
    return false;
  }
  data_attrs_i32 = xi;
  query_attrs_i32 = qi;
  return true;
}

void value_range(const Problem &p, int threads, double *lo, double *hi) {
  if (threads <= 0) threads = int(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  const int64_t nd = p.num_data * p.num_attrs, nq = p.num_queries * p.num_attrs;
  std::vector<double> los(size_t(threads), 1e300), his(size_t(threads), -1e300);
  auto scan = [&](int t) {
    double l = 1e300, h = -1e300;
    for (int pass = 0; pass < 2; ++pass) {
      const double *v = pass == 0 ? p.data_attrs : p.query_attrs;
      const int64_t n = pass == 0 ? nd : nq;
      for (int64_t i = n * t / threads, e = n * (t + 1) / threads; i < e; ++i) {
        l = v[i] < l ? v[i] : l;
        h = v[i] > h ? v[i] : h;
      }
    }
    los[size_t(t)] = l;
    his[size_t(t)] = h;
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(scan, t);
  scan(0);
  for (auto &th : pool) th.join();
  double l = 1e300, h = -1e300;
  for (int t = 0; t < threads; ++t) { l = los[size_t(t)] < l ? los[size_t(t)] : l; h = his[size_t(t)] > h ? his[size_t(t)] : h; }
  if (l > h) l = h = 0.0;
  *lo = l;
  *hi = h;
}

void synth_problem(const SynthSpec &s, Problem *out) {
  out->resize(s.n, s.q, s.a);
  int threads = int(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  int kmax = s.max_k < s.n ? s.max_k : int(s.n);
  int kmin = s.min_k < kmax ? s.min_k : kmax;
  if (kmin < 0) kmin = 0;
  auto fill = [&](int t) {
    int64_t b = s.n * t / threads, e = s.n * (t + 1) / threads;
    for (int64_t i = b; i < e; ++i) {
      out->data_labels[i] = int(synth_hash(s.seed, 1, uint64_t(i)) % uint64_t(s.num_labels));
      double *row = out->data_attrs + i * s.a;
      for (int j = 0; j < s.a; ++j)
        row[j] = synth_attr(synth_hash(s.seed, 0, uint64_t(i) * uint64_t(s.a) + uint64_t(j)), s.lo, s.hi);
    }
    b = s.q * t / threads; e = s.q * (t + 1) / threads;
    for (int64_t i = b; i < e; ++i) {
      out->query_k[i] = kmin + int(synth_hash(s.seed, 3, uint64_t(i)) % uint64_t(kmax - kmin + 1));
      double *row = out->query_attrs + i * s.a;
      for (int j = 0; j < s.a; ++j)
        row[j] = synth_attr(synth_hash(s.seed, 2, uint64_t(i) * uint64_t(s.a) + uint64_t(j)), s.lo, s.hi);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) pool.emplace_back(fill, t);
  for (auto &th : pool) th.join();
}

namespace {

// " %.6f" for one value.  Values that are exact six-decimal numbers (everything the
// generators produce) are printed with integer arithmetic; anything else goes through
// snprintf so the text is always what printf("%.6f") would give.
inline void append_fixed6(std::string &out, double v) {
  const double scaled = v * 1e6;
  if (std::fabs(scaled) < 9e15) {
    long long i = std::llround(scaled);
    if (static_cast<double>(i) / 1e6 == v) {
      char buf[32];
      int pos = 32;
      const bool neg = i < 0;
      unsigned long long u = neg ? static_cast<unsigned long long>(-i) : static_cast<unsigned long long>(i);
      for (int d = 0; d < 6; ++d) { buf[--pos] = char('0' + u % 10); u /= 10; }
      buf[--pos] = '.';
      do { buf[--pos] = char('0' + u % 10); u /= 10; } while (u != 0);
      if (neg && i != 0) buf[--pos] = '-';
      buf[--pos] = ' ';
      out.append(buf + pos, size_t(32 - pos));
      return;
    }
  }
  char tmp[64];
  int m = std::snprintf(tmp, sizeof tmp, " %.6f", v);
  out.append(tmp, size_t(m));
}

}  // namespace

void write_problem_text(const Problem &p, std::FILE *f) {
  std::fprintf(f, "%lld %lld %d\n", (long long)p.num_data, (long long)p.num_queries, p.num_attrs);
  int threads = int(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  if (threads > 32) threads = 32;
  const int64_t total = p.num_data + p.num_queries;
  // rows are formatted in parallel, block by block, and written in order
  const int64_t rows_per_thread = std::max<int64_t>(1, (8 << 20) / std::max(1, p.num_attrs * 12));
  const int64_t block = rows_per_thread * threads;
  std::vector<std::string> bufs;
  bufs.resize(static_cast<size_t>(threads));
  for (int64_t b0 = 0; b0 < total; b0 += block) {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
      pool.emplace_back([&, t] {
        std::string &line = bufs[size_t(t)];
        line.clear();
        const int64_t lo = b0 + t * rows_per_thread, hi = std::min(total, lo + rows_per_thread);
        for (int64_t r = lo; r < hi; ++r) {
          const double *row;
          if (r < p.num_data) {
            line += std::to_string(p.data_labels[r]);
            row = p.data_attrs + r * p.num_attrs;
          } else {
            line += "Q ";
            line += std::to_string(p.query_k[r - p.num_data]);
            row = p.query_attrs + (r - p.num_data) * p.num_attrs;
          }
          for (int j = 0; j < p.num_attrs; ++j) append_fixed6(line, row[j]);
          line += '\n';
        }
      });
    for (auto &th : pool) th.join();
    for (int t = 0; t < threads; ++t) std::fwrite(bufs[size_t(t)].data(), 1, bufs[size_t(t)].size(), f);
  }
}

}  // namespace dmlp
