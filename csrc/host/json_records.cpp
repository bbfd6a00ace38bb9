// Minimal, allocation-free scanner for newline-delimited DataInstance JSON records
// ({"numericalFeatures":[...],"discreteFeatures":[...],"target":x,"operation":"training"}).
// Replaces Jackson (utils/parsers/DataInstanceParser.scala:12-22) + the boxed-list to vector
// conversion (utils/parsers/dataStream/DataPointParser.scala:16-55) on the ingest path.
#include "omldm_host.h"

#include <cstdlib>
#include <cstring>

namespace {
const char* find_key(const char* b, const char* e, const char* key) {
  const size_t n = std::strlen(key);
  for (const char* p = b; p + n + 2 <= e; ++p)
    if (*p == '"' && std::memcmp(p + 1, key, n) == 0 && p[n + 1] == '"') {
      const char* q = p + n + 2;
      while (q < e && (*q == ' ' || *q == ':')) ++q;
      return q;
    }
  return nullptr;
}
// parses [a,b,c] of numbers into dst[pos...]; returns count or -1
int parse_array(const char* p, const char* e, float* dst, int pos, int cap) {
  if (!p || p >= e) return 0;
  if (std::strncmp(p, "null", 4) == 0) return 0;
  if (*p != '[') return -1;
  ++p;
  int n = 0;
  while (p < e) {
    while (p < e && (*p == ' ' || *p == ',')) ++p;
    if (p >= e) return -1;
    if (*p == ']') return n;
    char* end = nullptr;
    const float v = std::strtof(p, &end);
    if (end == p) return -1;
    if (pos + n < cap) dst[pos + n] = v;
    ++n;
    p = end;
  }
  return -1;
}
}  // namespace

extern "C" int omldm_parse_json_records(const char* text, size_t len, int d, float* dst, int ld, int max_rows,
                                        int* n_forecast, size_t* consumed) {
  int rows = 0, nf = 0;
  const char* p = text;
  const char* end = text + len;
  while (p < end && rows < max_rows) {
    const char* nl = static_cast<const char*>(std::memchr(p, '\n', end - p));
    const char* le = nl ? nl : end;
    const char* b = p;
    p = nl ? nl + 1 : end;
    while (b < le && (*b == ' ' || *b == '\r')) ++b;
    if (le - b < 2 || *b != '{') continue;   // blank line, "EOS", garbage: skipped silently
    float* row = dst + static_cast<size_t>(rows) * ld;
    std::memset(row, 0, sizeof(float) * ld);
    int n1 = parse_array(find_key(b, le, "numericalFeatures"), le, row, 0, d);
    if (n1 < 0) continue;
    int n2 = parse_array(find_key(b, le, "discreteFeatures"), le, row, n1, d);
    if (n2 < 0) continue;
    if (n1 + n2 != d) continue;              // wrong width for this stream
    const char* op = find_key(b, le, "operation");
    bool forecasting = false;
    if (op && op < le && *op == '"') {
      if (std::strncmp(op + 1, "forecasting", 11) == 0) forecasting = true;
      else if (std::strncmp(op + 1, "training", 8) != 0) continue;
    }
    float y = 0.f;
    const char* t = find_key(b, le, "target");
    if (t && std::strncmp(t, "null", 4) != 0) {
      char* e2 = nullptr;
      y = std::strtof(t, &e2);
      if (e2 == t) continue;
    }
    row[d] = forecasting ? 2.0f : 1.0f;
    row[d + 1] = y;
    if (forecasting) ++nf;
    ++rows;
  }
  if (n_forecast) *n_forecast = nf;
  if (consumed) *consumed = static_cast<size_t>(p - text);
  return rows;
}
