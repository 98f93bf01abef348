// nnom_blob.h -- host-side reader of data/kws_nnom_q7.bin (tools/convert_nnom_weights.py, made from the reference's
// firmware/src/ai/nnom/kws_nnom/weights.h).  Plain C++: shared by the library's loader and the host emulation harness.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

namespace edison {

struct KwsLayer {
  uint32_t h[9];          // kind, cout, kh, kw, cin, out_rshift, bias_lshift, relu, pool
  const int8_t *w, *b;    // into KwsBlob::bytes
};
struct KwsBlob {
  std::vector<unsigned char> bytes;
  KwsLayer L[5];
  int in_scale = 1, in_min = -128, in_max = 127;
};

// returns 0, or -4 (EDISON_E_DATA) / -6 (EDISON_E_MODEL) with a message in err
inline int read_kws_blob(const char *path, KwsBlob &out, char *err, size_t errlen) {
  FILE *fh = fopen(path, "rb");
  if (!fh) { snprintf(err, errlen, "cannot open %s", path); return -4; }
  fseek(fh, 0, SEEK_END);
  const long sz = ftell(fh);
  fseek(fh, 0, SEEK_SET);
  out.bytes.resize(sz > 0 ? (size_t)sz : 0);
  if (sz <= 0 || fread(out.bytes.data(), 1, (size_t)sz, fh) != (size_t)sz) { fclose(fh); snprintf(err, errlen, "%s: short read", path); return -4; }
  fclose(fh);
  const unsigned char *blob = out.bytes.data();
  if (sz < 16 || memcmp(blob, "EDNNQ7\0\0", 8)) { snprintf(err, errlen, "%s: bad header", path); return -4; }
  uint32_t ver, nl;
  memcpy(&ver, blob + 8, 4); memcpy(&nl, blob + 12, 4);
  if (ver != 1) { snprintf(err, errlen, "%s: version %u", path, ver); return -4; }
  // the graph this build has kernels for: kind, cout, kh, kw, cin, relu, pool
  static const uint32_t expect[5][7] = {{1, 16, 5, 5, 1, 1, 2}, {1, 32, 3, 3, 16, 1, 2}, {1, 64, 3, 3, 32, 1, 0}, {1, 32, 3, 3, 64, 1, 0}, {2, 10, 1, 1, 96, 0, 0}};
  if (nl != 5) { snprintf(err, errlen, "%s: %u layers, this build runs the 5-layer kws_conv graph", path, nl); return -6; }
  const unsigned char *p = blob + 16, *end = blob + sz;
  for (int i = 0; i < 5; i++) {
    if (p + 36 > end) { snprintf(err, errlen, "%s: truncated", path); return -4; }
    memcpy(out.L[i].h, p, 36); p += 36;
    const uint32_t *h = out.L[i].h;
    const uint32_t got[7] = {h[0], h[1], h[2], h[3], h[4], h[7], h[8]};
    if (memcmp(got, expect[i], sizeof got)) { snprintf(err, errlen, "%s: layer %d is not the kws_conv layer this build implements", path, i); return -6; }
    if (h[5] < 1 || h[5] > 23 || h[6] > 16) { snprintf(err, errlen, "%s: layer %d shifts out of range", path, i); return -4; }
    const size_t nw = (size_t)h[1] * h[2] * h[3] * h[4];
    if (p + nw + h[1] > end) { snprintf(err, errlen, "%s: truncated", path); return -4; }
    out.L[i].w = (const int8_t *)p; p += nw;
    out.L[i].b = (const int8_t *)p; p += h[1];
    p += (size_t)(-(p - blob)) & 3;
  }
  if (p + 24 > end) { snprintf(err, errlen, "%s: truncated trailer", path); return -4; }
  int32_t tq[3]; uint32_t sh[3];
  memcpy(tq, p, 12); memcpy(sh, p + 12, 12);
  if (sh[0] != 31 || sh[1] != 13 || sh[2] != 1 || tq[0] == 0) { snprintf(err, errlen, "%s: input shape/scale unsupported", path); return -6; }
  out.in_scale = tq[0]; out.in_min = tq[1]; out.in_max = tq[2];
  return 0;
}

}  // namespace edison
