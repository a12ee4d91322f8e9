// KITTI label text of the pseudo-boxes (host side of GenerateAnns, tools/PENet/dataloaders/my_loader.py:680):
//   cls 0 0 alpha xmin ymin xmax ymax h w l x y z ry\n
// The reference writes every float with Python's str(): the shortest decimal string that round-trips, fixed notation
// for 1e-4 <= |v| < 1e16 (always with a fractional part: "12.0"), else d.ddde+XX with at least two exponent digits.
// str(np.float64(v)) == repr(float(v)) == this routine (tests/test_cabi_symbols.py checks it against Python on 1e5
// values); std::to_chars yields the shortest digits, the layout below is CPython's float_repr_style = 'short'.
// ~4 k boxes x 8 doubles per 64-frame batch: 35 ms in Python, < 2 ms here.
#include <charconv>
#include <cmath>
#include "dfu_common.cuh"

// T = double: repr(float) / str(np.float64); T = float: str(np.float32) (numpy prints the shortest digits that round-trip
// in float32; its positional / scientific switch differs, see below)
template <typename T>
static char* put_repr(char* p, T v) {
  if (std::isnan(v)) { memcpy(p, "nan", 3); return p + 3; }
  if (std::isinf(v)) {
    if (v < 0) *p++ = '-';
    memcpy(p, "inf", 3);
    return p + 3;
  }
  char t[40];
  auto r = std::to_chars(t, t + sizeof(t), v, std::chars_format::scientific);   // [-]d[.ddd]e[+-]XX, shortest round trip
  const char* s = t;
  if (*s == '-') { *p++ = '-'; s++; }
  const char* e = s;
  while (e < r.ptr && *e != 'e') e++;
  char dig[24];
  int nd = 0;
  for (const char* q = s; q < e; q++)
    if (*q != '.') dig[nd++] = *q;
  int ex = 0;
  {
    const char* q = e + 1;
    bool neg = false;
    if (*q == '-') { neg = true; q++; } else if (*q == '+') q++;
    for (; q < r.ptr; q++) ex = ex * 10 + (*q - '0');
    if (neg) ex = -ex;
  }
  if (v == (T)0) { memcpy(p, "0.0", 3); return p + 3; }
  // Python's repr(float) chooses by the decimal exponent of the shortest digits (-4 <= ex < 16); numpy's str(np.float32)
  // by the VALUE (1e-4 <= |v| < 1e6: float32(0.0001) = 9.9999997e-05 prints as 1e-04)
  const double av = fabs((double)v);
  const bool positional = sizeof(T) == 8 ? (ex >= -4 && ex < 16) : (av >= 1e-4 && av < 1e6);
  if (positional) {
    if (ex < 0) {                                  // 0.000ddd
      *p++ = '0'; *p++ = '.';
      for (int i = 0; i < -ex - 1; i++) *p++ = '0';
      for (int i = 0; i < nd; i++) *p++ = dig[i];
    } else {
      for (int i = 0; i <= ex; i++) *p++ = i < nd ? dig[i] : '0';
      *p++ = '.';
      if (nd > ex + 1) for (int i = ex + 1; i < nd; i++) *p++ = dig[i];
      else *p++ = '0';
    }
  } else {                                         // d[.ddd]e+XX
    *p++ = dig[0];
    if (nd > 1) {
      *p++ = '.';
      for (int i = 1; i < nd; i++) *p++ = dig[i];
    }
    *p++ = 'e';
    *p++ = ex < 0 ? '-' : '+';
    int a = ex < 0 ? -ex : ex;
    char eb[8];
    int ne = 0;
    while (a > 0) { eb[ne++] = (char)('0' + a % 10); a /= 10; }
    while (ne < 2) eb[ne++] = '0';
    while (ne > 0) *p++ = eb[--ne];
  }
  return p;
}

// vals [n] -> out: the n strings, each followed by '\n'.  Returns the bytes written or -needed if cap is too small
// (48 bytes per value always suffice).
static char* put_repr_f64(char* p, double v) { return put_repr<double>(p, v); }

extern "C" int64_t dfu_repr_f64(const double* vals_host, int64_t n, char* out_host, int64_t cap) {
  if (!vals_host || !out_host || n < 0) return -1;
  if (cap < n * 48) return -(n * 48);
  char* p = out_host;
  for (int64_t i = 0; i < n; i++) {
    p = put_repr_f64(p, vals_host[i]);
    *p++ = '\n';
  }
  return (int64_t)(p - out_host);
}

// The four float32 tokens "xmin ymin xmax ymax" of every instance's 2-D box as str(np.float32) prints them
// (my_loader.py:680 formats box2d[i] values that are float32), NUL-terminated, off_host[i] = start of instance i: the b2txt /
// b2_off table of dfu_format_label_lines.  Returns the bytes written or -needed (100 bytes per instance always suffice).
extern "C" int64_t dfu_box2d_tokens(const float* b2_host, int64_t n_inst, char* out_host, int64_t cap, int32_t* off_host) {
  if (n_inst < 0 || (n_inst > 0 && (!b2_host || !out_host || !off_host))) return -1;
  if (cap < n_inst * 100) return -(n_inst * 100);
  char* p = out_host;
  for (int64_t i = 0; i < n_inst; i++) {
    off_host[i] = (int32_t)(p - out_host);
    for (int k = 0; k < 4; k++) {
      if (k) *p++ = ' ';
      p = put_repr<float>(p, b2_host[4 * i + k]);
    }
    *p++ = '\0';
  }
  return (int64_t)(p - out_host);
}

// All label lines of a batch.  boxes_host [n][DFU_BOX_DOUBLES] (records of dfu_fuse_instances_sync, sorted by view);
// names / b2txt: '\0'-separated strings (class names by class index; the four float32 2-D box tokens "xmin ymin xmax ymax"
// by GLOBAL instance index, formatted by the caller once per batch), name_off / b2_off their offsets.  Skipped records
// (field 20) produce no line.  lines_per_view_host [V] receives the line counts; the text goes to out_host, one line per
// kept record in record order.  Returns the bytes written or -needed.
extern "C" int64_t dfu_format_label_lines(const double* boxes_host, int64_t n, int V, const char* names, const int32_t* name_off,
                                          int n_names, const char* b2txt, const int32_t* b2_off, int64_t n_inst,
                                          char* out_host, int64_t cap, int64_t* lines_per_view_host) {
  if (n < 0 || V < 0 || (n > 0 && (!boxes_host || !names || !name_off || !b2txt || !b2_off || !out_host))) return -1;
  const int64_t need = n * 512;
  if (cap < need) return -need;
  if (lines_per_view_host) memset(lines_per_view_host, 0, sizeof(int64_t) * (size_t)V);
  char* p = out_host;
  for (int64_t i = 0; i < n; i++) {
    const double* r = boxes_host + i * DFU_BOX_DOUBLES;
    if (r[20] != 0.0) continue;
    const int v = (int)r[0], cls = (int)r[3];
    const int64_t g = (int64_t)r[1];
    if (v < 0 || v >= V || cls < 0 || cls >= n_names || g < 0 || g >= n_inst) return -1;
    const char* nm = names + name_off[cls];
    const size_t ln = strlen(nm);
    if (ln > 128) return -1;
    memcpy(p, nm, ln); p += ln;
    memcpy(p, " 0 0 ", 5); p += 5;
    p = put_repr_f64(p, r[4]);
    *p++ = ' ';
    const char* bt = b2txt + b2_off[g];
    const size_t lb = strlen(bt);
    if (lb > 128) return -1;
    memcpy(p, bt, lb); p += lb;
    for (int f = 5; f <= 11; f++) {
      *p++ = ' ';
      p = put_repr_f64(p, r[f]);
    }
    *p++ = '\n';
    if (lines_per_view_host) lines_per_view_host[v]++;
  }
  return (int64_t)(p - out_host);
}
