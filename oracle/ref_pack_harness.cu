// TEST INFRASTRUCTURE — not part of the product path.
//
// Tiny host-only driver around the reference's own packers, compiled from the
// reference header where it lies (/root/reference/mxmoe/kernels/src/include/quantize.cuh).
// Nothing from the reference is copied into this repo: the header is #included
// at build time and the binary lands in oracle/_ref/ (git-ignored).
//
// It exercises, on the CPU:
//   permute_scale    quantize.cuh:299-315
//   permute_weight   quantize.cuh:318-381
//   pack_weightonly  quantize.cuh:383-421
//   pack_wxax        quantize.cuh:424-475
//
// Protocol (binary, little endian) on stdin/stdout:
//   in : int32 op, N, K, M, qbits, sym, gsize ; then payload
//        op 0 (pack_weightonly): int32 q[N*K]          -> uint16 out[N*K/PACK]
//        op 1 (pack_wxax)      : int32 a[M*K], w[N*K]  -> uint16 a_out[M*K/PACK], w_out[N*K/PACK]
//        op 2 (permute_scale)  : uint16 s[N*K/g*(sym?1:2)] -> same count
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <tuple>
#include <vector>

#include "quantize.cuh"

template <typename T>
static void rd(T* p, size_t n) {
  if (fread(p, sizeof(T), n, stdin) != n) {
    fprintf(stderr, "short read\n");
    exit(2);
  }
}
template <typename T>
static void wr(const T* p, size_t n) {
  fwrite(p, sizeof(T), n, stdout);
}

int main() {
  int hdr[7];
  rd(hdr, 7);
  int op = hdr[0], N = hdr[1], K = hdr[2], M = hdr[3], qbits = hdr[4], sym = hdr[5], gsize = hdr[6];
  if (op == 0) {
    host_vector<int> q(size_t(N) * K);
    rd(q.data(), q.size());
    auto perm = permute_weight(q, N, K, qbits);
    auto out  = pack_weightonly(perm, N, K, qbits, sym);
    wr(reinterpret_cast<const uint16_t*>(out.data()), out.size());
  } else if (op == 1) {
    host_vector<int> a(size_t(M) * K), w(size_t(N) * K);
    rd(a.data(), a.size());
    rd(w.data(), w.size());
    if (qbits == 8) {
      auto [pa, pw] = pack_wxax<int8_t>(a, w, M, N, K, qbits);
      wr(reinterpret_cast<const uint16_t*>(pa.data()), pa.size());
      wr(reinterpret_cast<const uint16_t*>(pw.data()), pw.size());
    } else {
      auto [pa, pw] = pack_wxax<nv_precision::s4>(a, w, M, N, K, qbits);
      wr(reinterpret_cast<const uint16_t*>(pa.data()), pa.size());
      wr(reinterpret_cast<const uint16_t*>(pw.data()), pw.size());
    }
  } else if (op == 2) {
    size_t n = size_t(N) * (K / gsize) * (sym ? 1 : 2);
    host_vector<half> s(n);
    rd(reinterpret_cast<uint16_t*>(s.data()), n);
    auto out = permute_scale(s, N, K, gsize, qbits, sym);
    wr(reinterpret_cast<const uint16_t*>(out.data()), out.size());
  } else {
    return 3;
  }
  return 0;
}
