// LDA (Dirichlet) non-IID partition as a native host routine: bit-exact with
// /root/reference/data/noniid_partition.py:6-91 for the same numpy legacy MT19937 state.
//
// The reference draws from numpy's GLOBAL legacy generator (np.random.shuffle, np.random.dirichlet),
// so "the same seed" means: the same stream position before the call and the same position after it.
// This file therefore restates the pieces of numpy's legacy RandomState that the partition consumes -
// published algorithms, stated from their definitions, not copied code:
//   MT19937          Matsumoto & Nishimura 1998 (tempering 11 / 7,0x9d2c5680 / 15,0xefc60000 / 18)
//   double           53 bits from two outputs: (a >> 5) * 2^26 + (b >> 6), over 2^53
//   interval(max)    rejection under the smallest all-ones mask >= max, 32-bit draws while max < 2^32
//   shuffle          Fisher-Yates from the top: for i = n-1 .. 1: j = interval(i); swap(x[i], x[j])
//   gamma(shape<1)   Ahrens-Dieter GS as numpy's legacy code orders its draws (U, then the exponential V)
//   gamma(shape>1)   Marsaglia-Tsang with the legacy polar gaussian (one value cached)
//   dirichlet        K gammas normalised by their sum
// and Python's random.shuffle (CPython: getrandbits(k) = top k bits of one 32-bit output, rejection
// until < n), which the reference's train/test split uses on every client's id list
// (relational_table_preprocessor.py:83).  tests/test_host_logic.py checks all of it against numpy and
// the `random` module, stream position included.
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/splitnn.h"

namespace {

struct Mt {
  uint32_t* key;   // [624]
  int* pos;
  uint32_t next() {
    if (*pos >= 624) {
      uint32_t* mt = key;
      for (int i = 0; i < 624; ++i) {
        const uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
        mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
      }
      *pos = 0;
    }
    uint32_t y = key[(*pos)++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
  }
  double dbl() {
    const uint32_t a = next() >> 5, b = next() >> 6;
    return (a * 67108864.0 + b) / 9007199254740992.0;
  }
  uint64_t interval(uint64_t max) {
    if (max == 0) return 0;
    uint64_t mask = max;
    mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16; mask |= mask >> 32;
    uint64_t v;
    if (max <= 0xffffffffull) {
      while ((v = (next() & mask)) > max) {}
    } else {
      for (;;) {
        const uint64_t hi = next(), lo = next();
        v = ((hi << 32) | lo) & mask;
        if (v <= max) break;
      }
    }
    return v;
  }
};

struct Legacy {
  Mt mt;
  int* has_gauss;
  double* gauss;
  double gaussian() {
    if (*has_gauss) {
      const double t = *gauss;
      *has_gauss = 0;
      *gauss = 0.0;
      return t;
    }
    double x1, x2, r2;
    do {
      x1 = 2.0 * mt.dbl() - 1.0;
      x2 = 2.0 * mt.dbl() - 1.0;
      r2 = x1 * x1 + x2 * x2;
    } while (r2 >= 1.0 || r2 == 0.0);
    const double f = sqrt(-2.0 * log(r2) / r2);
    *gauss = f * x1;
    *has_gauss = 1;
    return f * x2;
  }
  double exponential() { return -log(1.0 - mt.dbl()); }
  double gamma(double shape) {
    if (shape == 1.0) return exponential();
    if (shape == 0.0) return 0.0;
    if (shape < 1.0) {
      for (;;) {
        const double U = mt.dbl();
        const double V = exponential();
        if (U <= 1.0 - shape) {
          const double X = pow(U, 1.0 / shape);
          if (X <= V) return X;
        } else {
          const double Y = -log((1.0 - U) / shape);
          const double X = pow(1.0 - shape + shape * Y, 1.0 / shape);
          if (X <= V + Y) return X;
        }
      }
    }
    const double b = shape - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * b);
    for (;;) {
      double X, V;
      do {
        X = gaussian();
        V = 1.0 + c * X;
      } while (V <= 0.0);
      V = V * V * V;
      const double U = mt.dbl();
      if (U < 1.0 - 0.0331 * (X * X) * (X * X)) return b * V;
      if (log(U) < 0.5 * X * X + b * (1.0 - V + log(V))) return b * V;
    }
  }
};

// ndarray.sum() of a contiguous float64 vector adds in numpy's pairwise order, not left to right:
// under 8 values a plain loop; up to 128 eight running sums combined as a tree, the tail added after;
// above that the halves (the left one a multiple of 8) summed recursively.
double pairwise_sum(const double* a, int64_t n) {
  if (n < 8) {
    double res = 0.0;
    for (int64_t i = 0; i < n; ++i) res += a[i];
    return res;
  }
  if (n <= 128) {
    double r[8];
    for (int q = 0; q < 8; ++q) r[q] = a[q];
    int64_t i = 8;
    for (; i < n - (n % 8); i += 8)
      for (int q = 0; q < 8; ++q) r[q] += a[i + q];
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; ++i) res += a[i];
    return res;
  }
  int64_t n2 = n / 2;
  n2 -= n2 % 8;
  return pairwise_sum(a, n2) + pairwise_sum(a + n2, n - n2);
}

template <typename T>
void shuffle(Mt& mt, T* x, int64_t n) {
  for (int64_t i = n - 1; i >= 1; --i) {
    const int64_t j = (int64_t)mt.interval((uint64_t)i);
    std::swap(x[i], x[j]);
  }
}

}  // namespace

extern "C" {

int sn_lda_partition(const int64_t* labels, int64_t n, int client_num, int classes, double alpha, uint32_t* mt_key624,
                     int* mt_pos, int* has_gauss, double* cached_gauss, int64_t* out_idx, int64_t* out_offsets) {
  if (!labels || n <= 0 || client_num <= 0 || classes <= 0 || !mt_key624 || !mt_pos || !has_gauss || !cached_gauss || !out_idx ||
      !out_offsets || *mt_pos < 0 || *mt_pos > 624)
    return 2;
  Legacy rng{Mt{mt_key624, mt_pos}, has_gauss, cached_gauss};
  // class k's sample ids in ascending order: np.where(label_list == k)[0]  (noniid_partition.py:66)
  std::vector<std::vector<int64_t>> by_class(classes);
  for (int64_t i = 0; i < n; ++i)
    if (labels[i] >= 0 && labels[i] < classes) by_class[labels[i]].push_back(i);
  const double cap = (double)n / (double)client_num;        // N / client_num (:83)
  std::vector<std::vector<int64_t>> shards;
  std::vector<double> prop(client_num);
  int64_t smallest = 0;
  while (smallest < 10) {                                     // min-size guarantee, restart from scratch (:43-44)
    shards.assign(client_num, {});
    for (int k = 0; k < classes; ++k) {
      std::vector<int64_t> idx = by_class[k];
      shuffle(rng.mt, idx.data(), (int64_t)idx.size());       // np.random.shuffle(idx_k) (:77)
      double acc = 0.0;                                       // np.random.dirichlet(repeat(alpha, K)) (:80)
      for (int j = 0; j < client_num; ++j) {
        prop[j] = rng.gamma(alpha);
        acc += prop[j];
      }
      const double inv = 1.0 / acc;
      for (int j = 0; j < client_num; ++j) prop[j] *= inv;
      for (int j = 0; j < client_num; ++j)                    // p * (len(idx_j) < N/K), renormalised (:83-84)
        prop[j] = prop[j] * (((double)shards[j].size() < cap) ? 1.0 : 0.0);
      const double sum = pairwise_sum(prop.data(), client_num);
      double run = 0.0;                                       // (cumsum(p) * len(idx_k)).astype(int)[:-1] (:85)
      int64_t begin = 0;
      const int64_t len = (int64_t)idx.size();
      for (int j = 0; j < client_num; ++j) {
        run += prop[j] / sum;
        int64_t end = (j == client_num - 1) ? len : (int64_t)(run * (double)len);
        end = std::max<int64_t>(begin, std::min<int64_t>(end, len));      // np.split clips to the array
        shards[j].insert(shards[j].end(), idx.begin() + begin, idx.begin() + end);
        begin = end;
      }
    }
    smallest = (int64_t)shards[0].size();
    for (auto& s : shards) smallest = std::min<int64_t>(smallest, (int64_t)s.size());
  }
  int64_t o = 0;
  for (int i = 0; i < client_num; ++i) {                      // np.random.shuffle(idx_batch[i]) (:69-71)
    shuffle(rng.mt, shards[i].data(), (int64_t)shards[i].size());
    out_offsets[i] = o;
    memcpy(out_idx + o, shards[i].data(), shards[i].size() * sizeof(int64_t));
    o += (int64_t)shards[i].size();
  }
  out_offsets[client_num] = o;
  return 0;
}

int sn_py_random_shuffle(int64_t* ids, int64_t n, uint32_t* mt_key624, int* mt_pos) {
  if (!ids || n < 0 || !mt_key624 || !mt_pos || *mt_pos < 0 || *mt_pos > 624) return 2;
  Mt mt{mt_key624, mt_pos};
  for (int64_t i = n - 1; i >= 1; --i) {                      // random.shuffle: j = randbelow(i + 1)
    const uint64_t bound = (uint64_t)i + 1;
    int k = 0;
    while ((bound >> k) != 0) ++k;                            // bound.bit_length()
    uint64_t r;
    do {
      if (k <= 32) {
        r = mt.next() >> (32 - k);
      } else {                                                // getrandbits(k > 32): 32-bit words, least significant first
        const uint64_t lo = mt.next();
        const uint64_t hi = mt.next() >> (64 - k);
        r = (hi << 32) | lo;
      }
    } while (r >= bound);
    std::swap(ids[i], ids[(int64_t)r]);
  }
  return 0;
}

}  // extern "C"
