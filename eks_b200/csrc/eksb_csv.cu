// eksb_csv.cu -- host-side I/O for the on-disk format either side of the hot path:
// DeepLabCut / Lightning-Pose prediction CSVs -- 3 header rows (scorer / bodyparts / coords), one index
// column, then a dense numeric body.  The reference reads them with
// pd.read_csv(path, header=[0,1,2], index_col=0) (scripts/multicam_example.py:85) and writes results with
// DataFrame.to_csv (:139,:157).  Here the numeric body is parsed by a multi-threaded std::from_chars
// scanner straight into a caller-provided (rows, cols) float64 array (pinned host memory when the caller
// wants to hand it to cudaMemcpyAsync), and written back with the shortest round-trip digits laid out
// the way Python's float repr does (what to_csv emits), empty field for NaN.
// No device code in this file; it lives in libeksb so that one C ABI covers the path end to end.
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "eksb_common.cuh"

namespace eksb {

static bool read_file(const char* path, std::string& buf) {
  FILE* f = fopen(path, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  buf.resize(n > 0 ? (size_t)n : 0);
  const size_t got = n > 0 ? fread(&buf[0], 1, (size_t)n, f) : 0;
  fclose(f);
  return got == buf.size();
}

// offsets of the starts of all non-empty lines after `skip` header lines; returns header end offset
static size_t index_lines(const std::string& b, int skip, std::vector<size_t>& starts) {
  size_t pos = 0;
  for (int i = 0; i < skip && pos < b.size(); ++i) {
    const size_t e = b.find('\n', pos);
    pos = (e == std::string::npos) ? b.size() : e + 1;
  }
  const size_t body = pos;
  while (pos < b.size()) {
    size_t e = b.find('\n', pos);
    if (e == std::string::npos) e = b.size();
    if (e > pos && !(e == pos + 1 && b[pos] == '\r')) starts.push_back(pos);
    pos = e + 1;
  }
  return body;
}

static int count_fields(const std::string& b, size_t start) {
  int n = 1;
  for (size_t i = start; i < b.size() && b[i] != '\n'; ++i) n += (b[i] == ',');
  return n;
}

// parse one line: first field -> index (integer, may be absent), remaining n_cols fields -> doubles
static bool parse_line(const char* p, const char* end, long* index, double* out, int n_cols) {
  const char* q = p;
  while (q < end && *q != ',') ++q;
  if (index) {
    long v = 0;
    auto r = std::from_chars(p, q, v);
    *index = (r.ec == std::errc()) ? v : -1;
  }
  p = q < end ? q + 1 : end;
  for (int j = 0; j < n_cols; ++j) {
    q = p;
    while (q < end && *q != ',' && *q != '\r') ++q;
    double v = NAN;
    if (q > p) {
      const char* s = p;
      if (*s == '+') ++s;
      auto r = std::from_chars(s, q, v);
      if (r.ec != std::errc()) {
        if (r.ec == std::errc::result_out_of_range) v = (*s == '-') ? -INFINITY : INFINITY;
        else if ((q - s) >= 3 && (s[0] == 'n' || s[0] == 'N')) v = NAN;
        else return false;
      }
    }
    out[j] = v;
    if (q >= end || *q == '\r') {
      if (j != n_cols - 1) return false;
      break;
    }
    p = q + 1;
  }
  return true;
}

}  // namespace eksb

using namespace eksb;

extern "C" int eksb_csv_dlc_shape_host(const char* path, long* n_rows, int* n_cols) {
  if (!path || !n_rows || !n_cols) { set_error("csv_dlc_shape: null argument"); return EKSB_EARG; }
  std::string buf;
  if (!read_file(path, buf)) { set_error("csv_dlc_shape: cannot read %s", path); return EKSB_EARG; }
  std::vector<size_t> starts;
  index_lines(buf, 3, starts);
  *n_rows = (long)starts.size();
  *n_cols = count_fields(buf, 0) - 1;
  return EKSB_OK;
}

extern "C" int eksb_csv_dlc_read_host(const char* path, double* data_host, long n_rows, int n_cols, long* index_host,
                                      char* header_host, size_t header_cap, int n_threads) {
  if (!path || !data_host || n_rows < 0 || n_cols <= 0) { set_error("csv_dlc_read: bad argument"); return EKSB_EARG; }
  std::string buf;
  if (!read_file(path, buf)) { set_error("csv_dlc_read: cannot read %s", path); return EKSB_EARG; }
  std::vector<size_t> starts;
  const size_t body = index_lines(buf, 3, starts);
  if ((long)starts.size() != n_rows) { set_error("csv_dlc_read: file has %zu rows, caller expects %ld", starts.size(), n_rows); return EKSB_EARG; }
  if (header_host && header_cap) {
    const size_t n = body < header_cap - 1 ? body : header_cap - 1;
    memcpy(header_host, buf.data(), n);
    header_host[n] = 0;
  }
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if (nt > 64) nt = 64;
  if ((long)nt > n_rows / 256 + 1) nt = (int)(n_rows / 256 + 1);
  std::vector<int> bad(nt, 0);
  auto work = [&](int w) {
    const long r0 = n_rows * w / nt, r1 = n_rows * (w + 1) / nt;
    for (long r = r0; r < r1; ++r) {
      const char* p = buf.data() + starts[r];
      const char* e = (const char*)memchr(p, '\n', buf.size() - starts[r]);
      if (!e) e = buf.data() + buf.size();
      if (!parse_line(p, e, index_host ? index_host + r : nullptr, data_host + r * n_cols, n_cols)) { bad[w] = 1; return; }
    }
  };
  std::vector<std::thread> th;
  for (int w = 1; w < nt; ++w) th.emplace_back(work, w);
  work(0);
  for (auto& t : th) t.join();
  for (int w = 0; w < nt; ++w)
    if (bad[w]) { set_error("csv_dlc_read: malformed numeric field in %s", path); return EKSB_EARG; }
  return EKSB_OK;
}

// shortest round-trip decimal, laid out like Python's repr(float) (what DataFrame.to_csv writes)
static int format_repr(double v, char* out) {
  if (std::isnan(v)) return 0;                               // to_csv: empty field
  if (std::isinf(v)) return sprintf(out, v < 0 ? "-inf" : "inf");
  char sci[40];
  auto r = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);
  *r.ptr = 0;
  // sci = [-]d[.ddd]e[+-]xx
  char* p = sci;
  char* o = out;
  if (*p == '-') *o++ = *p++;
  char digits[24];
  int nd = 0;
  for (; *p && *p != 'e'; ++p)
    if (*p != '.') digits[nd++] = *p;
  const int exp10 = atoi(p + 1);
  const int decpt = exp10 + 1;                              // position of the decimal point after digit 0
  if (decpt > 16 || decpt < -3) {                           // repr's scientific branch: d.ddde+XX
    *o++ = digits[0];
    if (nd > 1) { *o++ = '.'; for (int i = 1; i < nd; ++i) *o++ = digits[i]; }
    o += sprintf(o, "e%c%02d", exp10 < 0 ? '-' : '+', exp10 < 0 ? -exp10 : exp10);
  } else if (decpt <= 0) {
    *o++ = '0'; *o++ = '.';
    for (int i = 0; i < -decpt; ++i) *o++ = '0';
    for (int i = 0; i < nd; ++i) *o++ = digits[i];
  } else {
    for (int i = 0; i < decpt; ++i) *o++ = i < nd ? digits[i] : '0';
    *o++ = '.';
    if (nd > decpt) for (int i = decpt; i < nd; ++i) *o++ = digits[i];
    else *o++ = '0';
  }
  *o = 0;
  return (int)(o - out);
}

extern "C" int eksb_csv_dlc_write_host(const char* path, const char* header_host, const long* index_host,
                                       const double* data_host, long n_rows, int n_cols) {
  if (!path || !header_host || !data_host || n_rows < 0 || n_cols <= 0) { set_error("csv_dlc_write: bad argument"); return EKSB_EARG; }
  FILE* f = fopen(path, "wb");
  if (!f) { set_error("csv_dlc_write: cannot open %s", path); return EKSB_EARG; }
  fputs(header_host, f);
  const size_t hl = strlen(header_host);
  if (hl && header_host[hl - 1] != '\n') fputc('\n', f);
  std::string line;
  char num[48];
  for (long r = 0; r < n_rows; ++r) {
    line.clear();
    line += std::to_string(index_host ? index_host[r] : r);
    for (int j = 0; j < n_cols; ++j) {
      line += ',';
      const int n = format_repr(data_host[r * n_cols + j], num);
      line.append(num, (size_t)n);
    }
    line += '\n';
    fwrite(line.data(), 1, line.size(), f);
  }
  fclose(f);
  return EKSB_OK;
}
