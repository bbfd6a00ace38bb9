// Allocation-free scanner for newline-delimited DataInstance JSON records
// ({"numericalFeatures":[...],"discreteFeatures":[...],"target":x,"operation":"training"}).
// Replaces Jackson (utils/parsers/DataInstanceParser.scala:12-22) + the boxed-list to vector
// conversion (utils/parsers/dataStream/DataPointParser.scala:16-55) on the ingest path.
//
// One pass per line: keys are dispatched as they are met, unknown values are skipped structurally;
// numbers go through a 19-digit fixed-point fast path (correctly rounded double for mantissas < 2^53 and
// |exp10| <= 22, within a few double ulps otherwise - the rows are float32; strtod beyond 1e+-300).  The multi-threaded entry point splits the text at line boundaries.
#include "omldm_host.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

const double kPow10[] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                         1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// Parses a JSON number at p (p < e).  Returns the position after it, or nullptr if there is no number.
inline const char* parse_number(const char* p, const char* e, double* out) {
  const char* s = p;
  bool neg = false;
  if (p < e && (*p == '-' || *p == '+')) { neg = *p == '-'; ++p; }
  uint64_t mant = 0;
  int digits = 0, dropped = 0, frac = 0;
  bool any = false;
  while (p < e && *p >= '0' && *p <= '9') {
    any = true;
    if (digits < 19) { mant = mant * 10 + static_cast<uint64_t>(*p - '0'); if (mant) ++digits; }
    else ++dropped;
    ++p;
  }
  if (p < e && *p == '.') {
    ++p;
    while (p < e && *p >= '0' && *p <= '9') {
      any = true;
      if (digits < 19) { mant = mant * 10 + static_cast<uint64_t>(*p - '0'); if (mant) ++digits; ++frac; }
      ++p;
    }
  }
  if (!any) return nullptr;
  int ex = 0;
  if (p < e && (*p == 'e' || *p == 'E')) {
    const char* q = p + 1;
    bool eneg = false;
    if (q < e && (*q == '-' || *q == '+')) { eneg = *q == '-'; ++q; }
    if (q < e && *q >= '0' && *q <= '9') {
      int v = 0;
      while (q < e && *q >= '0' && *q <= '9') { if (v < 100000) v = v * 10 + (*q - '0'); ++q; }
      ex = eneg ? -v : v;
      p = q;
    }
  }
  const int e10 = ex + dropped - frac;
  if (e10 >= -300 && e10 <= 300) {
    // mant < 2^53 and |e10| <= 22: one exact operand pair, correctly rounded.  Otherwise each step adds at
    // most half a double ulp - far below the float32 precision the rows are stored in.
    double v = static_cast<double>(mant);
    int k = e10;
    while (k < 0) { const int s10 = k < -22 ? 22 : -k; v /= kPow10[s10]; k += s10; }
    while (k > 0) { const int s10 = k > 22 ? 22 : k; v *= kPow10[s10]; k -= s10; }
    *out = neg ? -v : v;
    return p;
  }
  char buf[64];                                    // rare: exponents beyond +-300
  const size_t n = static_cast<size_t>(p - s);
  if (n >= sizeof(buf)) { *out = std::strtod(std::string(s, n).c_str(), nullptr); return p; }
  std::memcpy(buf, s, n);
  buf[n] = 0;
  *out = std::strtod(buf, nullptr);
  return p;
}

inline const char* skip_ws(const char* p, const char* e) {
  while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) ++p;
  return p;
}

// p at the opening quote; returns the position after the closing quote (escapes honoured) or nullptr
inline const char* skip_string(const char* p, const char* e) {
  ++p;
  while (p < e) {
    if (*p == '\\') { p += 2; continue; }
    if (*p == '"') return p + 1;
    ++p;
  }
  return nullptr;
}

// skips any JSON value; returns the position after it or nullptr
const char* skip_value(const char* p, const char* e) {
  p = skip_ws(p, e);
  if (p >= e) return nullptr;
  if (*p == '"') return skip_string(p, e);
  if (*p == '[' || *p == '{') {
    int depth = 0;
    while (p < e) {
      if (*p == '"') { p = skip_string(p, e); if (!p) return nullptr; continue; }
      if (*p == '[' || *p == '{') ++depth;
      else if (*p == ']' || *p == '}') { if (--depth == 0) return p + 1; }
      ++p;
    }
    return nullptr;
  }
  while (p < e && *p != ',' && *p != '}' && *p != ']') ++p;   // number / true / false / null
  return p;
}

// [a,b,c] or null -> dst[0..cap); returns the position after the value and the count in *n (-1: malformed)
const char* parse_array(const char* p, const char* e, float* dst, int cap, int* n) {
  p = skip_ws(p, e);
  *n = 0;
  if (e - p >= 4 && std::memcmp(p, "null", 4) == 0) return p + 4;
  if (p >= e || *p != '[') { *n = -1; return nullptr; }
  ++p;
  int k = 0;
  while (p < e) {
    while (p < e && (*p == ' ' || *p == ',' || *p == '\t')) ++p;
    if (p >= e) break;
    if (*p == ']') { *n = k; return p + 1; }
    double v;
    const char* q = parse_number(p, e, &v);
    if (!q) break;
    if (k < cap) dst[k] = static_cast<float>(v);
    ++k;
    p = q;
  }
  *n = -1;
  return nullptr;
}

// One record line [b, le) -> row.  Returns 0 skipped, 1 training, 2 forecasting.
int parse_line(const char* b, const char* le, int d, float* row, int ld, float* tmp) {
  b = skip_ws(b, le);
  if (le - b < 2 || *b != '{') return 0;           // blank line, "EOS", garbage: skipped silently
  std::memset(row, 0, sizeof(float) * static_cast<size_t>(ld));
  int n1 = 0, n2 = 0;
  bool forecasting = false;
  float y = 0.f;
  const char* p = b + 1;
  while (p < le) {
    while (p < le && *p != '"' && *p != '}') ++p;
    if (p >= le || *p == '}') break;
    const char* k0 = p + 1;
    const char* k1 = skip_string(p, le);
    if (!k1) return 0;
    const size_t kl = static_cast<size_t>(k1 - 1 - k0);
    p = skip_ws(k1, le);
    if (p >= le || *p != ':') return 0;
    p = skip_ws(p + 1, le);
    if (kl == 17 && std::memcmp(k0, "numericalFeatures", 17) == 0) {
      p = parse_array(p, le, row, d, &n1);
      if (!p) return 0;
    } else if (kl == 16 && std::memcmp(k0, "discreteFeatures", 16) == 0) {
      p = parse_array(p, le, tmp, d, &n2);
      if (!p) return 0;
    } else if (kl == 6 && std::memcmp(k0, "target", 6) == 0) {
      if (le - p >= 4 && std::memcmp(p, "null", 4) == 0) {
        p += 4;
      } else {
        double v;
        const char* q = parse_number(p, le, &v);
        if (!q) return 0;
        y = static_cast<float>(v);
        p = q;
      }
    } else if (kl == 9 && std::memcmp(k0, "operation", 9) == 0) {
      if (p < le && *p == '"') {
        const char* q = skip_string(p, le);
        if (!q) return 0;
        const size_t vl = static_cast<size_t>(q - 1 - (p + 1));
        if (vl == 11 && std::memcmp(p + 1, "forecasting", 11) == 0) forecasting = true;
        else if (!(vl == 8 && std::memcmp(p + 1, "training", 8) == 0)) return 0;   // unknown operation: dropped
        p = q;
      } else {
        p = skip_value(p, le);
        if (!p) return 0;
      }
    } else {
      p = skip_value(p, le);
      if (!p) return 0;
    }
  }
  if (n1 + n2 != d) return 0;                      // wrong width for this stream
  for (int j = 0; j < n2; ++j) row[n1 + j] = tmp[j];
  row[d] = forecasting ? 2.0f : 1.0f;
  row[d + 1] = y;
  return forecasting ? 2 : 1;
}

}  // namespace

extern "C" int omldm_parse_json_records(const char* text, size_t len, int d, float* dst, int ld, int max_rows,
                                        int* n_forecast, size_t* consumed) {
  int rows = 0, nf = 0;
  const char* p = text;
  const char* end = text + len;
  std::vector<float> tmp(static_cast<size_t>(d > 0 ? d : 1));
  while (p < end && rows < max_rows) {
    const char* nl = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(end - p)));
    const char* le = nl ? nl : end;
    const char* b = p;
    p = nl ? nl + 1 : end;
    const int r = parse_line(b, le, d, dst + static_cast<size_t>(rows) * ld, ld, tmp.data());
    if (r == 0) continue;
    if (r == 2) ++nf;
    ++rows;
  }
  if (n_forecast) *n_forecast = nf;
  if (consumed) *consumed = static_cast<size_t>(p - text);
  return rows;
}

// Multi-threaded variant: at most max_rows LINES are taken from the text (so at most max_rows rows come out);
// each thread parses a contiguous range of lines into the rows of the same numbers, skipped lines are squeezed
// out afterwards.  Row order = line order.
extern "C" int omldm_parse_json_records_mt(const char* text, size_t len, int d, float* dst, int ld, int max_rows,
                                           int* n_forecast, size_t* consumed, int n_threads) {
  if (n_threads <= 1 || len < (1u << 16))
    return omldm_parse_json_records(text, len, d, dst, ld, max_rows, n_forecast, consumed);
  std::vector<const char*> starts;
  starts.reserve(static_cast<size_t>(std::min<size_t>(static_cast<size_t>(max_rows), len / 16 + 1)) + 1);
  const char* p = text;
  const char* end = text + len;
  while (p < end && static_cast<int>(starts.size()) < max_rows) {
    starts.push_back(p);
    const char* nl = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(end - p)));
    p = nl ? nl + 1 : end;
  }
  starts.push_back(p);                               // one past the last line taken
  const int lines = static_cast<int>(starts.size()) - 1;
  const int T = std::min(n_threads, std::max(1, lines / 256));
  std::vector<unsigned char> kind(static_cast<size_t>(lines), 0);
  auto work = [&](int t) {
    const int lo = static_cast<int>(static_cast<long long>(lines) * t / T);
    const int hi = static_cast<int>(static_cast<long long>(lines) * (t + 1) / T);
    std::vector<float> tmp(static_cast<size_t>(d > 0 ? d : 1));
    for (int i = lo; i < hi; ++i) {
      const char* b = starts[static_cast<size_t>(i)];
      const char* le = starts[static_cast<size_t>(i) + 1];
      if (le > b && le[-1] == '\n') --le;
      kind[static_cast<size_t>(i)] =
          static_cast<unsigned char>(parse_line(b, le, d, dst + static_cast<size_t>(i) * ld, ld, tmp.data()));
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
  int rows = 0, nf = 0;
  for (int i = 0; i < lines; ++i) {
    if (!kind[static_cast<size_t>(i)]) continue;
    if (rows != i)
      std::memcpy(dst + static_cast<size_t>(rows) * ld, dst + static_cast<size_t>(i) * ld, sizeof(float) * static_cast<size_t>(ld));
    if (kind[static_cast<size_t>(i)] == 2) ++nf;
    ++rows;
  }
  if (n_forecast) *n_forecast = nf;
  if (consumed) *consumed = static_cast<size_t>(p - text);
  return rows;
}
