// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of Apache Spark MLlib 3.1.2 tree training (the arithmetic behind
// pyspark.ml.regression.DecisionTreeRegressor / RandomForestRegressor as called from the reference
// notebooks `ML 06 - Decision Trees.py:73-75,118,126` and
// `ML 07 - Random Forests and Hyperparameter Tuning.py:41,74-77,107-116`).
//
// The algorithm lives in a third-party dependency that is NOT vendored under /root/reference:
//   org.apache.spark:spark-mllib_2.12:3.1.2 (+ commons-math3 3.4.1, java.util.Random, Scala 2.12 collections)
// (pin inferred from `Includes/Class-Utility-Methods.py:93-94`, DBR 9.1).  Every function below cites the
// upstream class it restates and the SURVEY.md Appendix A paragraph that is its spec.
//
// PARITY STATUS: pinned only by the known-answer tests of SURVEY.md Appendix B (upstream unit-test values and
// pyspark doctest values); the reference repo itself stores no outputs for this path, and no JVM exists in this
// image, so agreement with a live MLlib run is "parity unpinned" (see DESIGN.md).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

thread_local std::string g_err;

// ----------------------------------------------------------------------------------------------
// A.5 RNG primitives
// ----------------------------------------------------------------------------------------------

// java.util.Random (48-bit LCG).  Spec: Appendix A.5; KAT Appendix B "java.util.Random".
struct JavaRandom {
  uint64_t s;
  explicit JavaRandom(int64_t seed) { s = ((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
  int32_t next(int bits) {
    s = (s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
    return (int32_t)((int64_t)s >> (48 - bits));
  }
  int32_t nextInt() { return next(32); }
  int64_t nextLong() { return (int64_t)((uint64_t)(int64_t)next(32) << 32) + (int64_t)next(32); }
  double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
};

// scala.util.hashing.MurmurHash3.bytesHash (== MurmurHash3_x86_32).  Appendix A.5.
inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
uint32_t murmur3_bytes(const uint8_t* data, int len, uint32_t seed) {
  uint32_t h = seed;
  int i = 0;
  for (; i + 4 <= len; i += 4) {
    uint32_t k = (uint32_t)data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16) | ((uint32_t)data[i + 3] << 24);
    k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u;
    h ^= k; h = rotl32(h, 13); h = h * 5 + 0xe6546b64u;
  }
  uint32_t k = 0;
  int rem = len - i;
  if (rem == 3) k ^= (uint32_t)data[i + 2] << 16;
  if (rem >= 2) k ^= (uint32_t)data[i + 1] << 8;
  if (rem >= 1) { k ^= (uint32_t)data[i]; k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u; h ^= k; }
  h ^= (uint32_t)len;
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}

// org.apache.spark.util.random.XORShiftRandom.  Appendix A.5; KAT Appendix B "XORShiftRandom".
struct XORShiftRandom {
  uint64_t s;
  static uint64_t hashSeed(int64_t seed) {
    uint8_t b[8];
    for (int i = 0; i < 8; i++) b[i] = (uint8_t)((uint64_t)seed >> (56 - 8 * i));  // ByteBuffer.putLong: big endian
    uint32_t lo = murmur3_bytes(b, 8, 0x3c074a61u);
    uint32_t hi = murmur3_bytes(b, 8, lo);
    return ((uint64_t)hi << 32) | (uint64_t)lo;
  }
  explicit XORShiftRandom(int64_t init) : s(hashSeed(init)) {}
  int32_t next(int bits) {
    uint64_t n = s ^ (s << 21);
    n ^= n >> 35;
    n ^= n << 4;
    s = n;
    return (int32_t)(n & ((1ULL << bits) - 1));
  }
  int32_t nextInt() { return next(32); }
  double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
};

// org.apache.commons.math3.random.Well19937c (AbstractWell seeding).  Appendix A.4.
struct Well19937c {
  static constexpr int R = 624, M1 = 70, M2 = 179, M3 = 449;
  uint32_t v[R];
  int index;
  explicit Well19937c(int64_t seed) { setSeed(seed); }
  void setSeed(int64_t seed) {
    uint32_t s2[2] = {(uint32_t)((uint64_t)seed >> 32), (uint32_t)((uint64_t)seed & 0xffffffffULL)};
    v[0] = s2[0]; v[1] = s2[1];
    for (int i = 2; i < R; i++) {
      int64_t l = (int64_t)(int32_t)v[i - 2];  // sign-extended int
      v[i] = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
    }
    index = 0;
  }
  uint32_t next(int bits) {
    const int iRm1 = (index + R - 1) % R, iRm2 = (index + R - 2) % R;
    const uint32_t v0 = v[index], vM1 = v[(index + M1) % R], vM2 = v[(index + M2) % R], vM3 = v[(index + M3) % R];
    const uint32_t z0 = (0x80000000u & v[iRm1]) ^ (0x7FFFFFFFu & v[iRm2]);
    const uint32_t z1 = (v0 ^ (v0 << 25)) ^ (vM1 ^ (vM1 >> 27));
    const uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
    const uint32_t z3 = z1 ^ z2;
    uint32_t z4 = z0 ^ (z1 ^ (z1 << 9)) ^ (z2 ^ (z2 << 21)) ^ (z3 ^ (z3 >> 21));
    v[index] = z3;
    v[iRm1] = z4;
    v[iRm2] &= 0x80000000u;
    index = iRm1;
    z4 ^= (z4 << 7) & 0xe46e1700u;
    z4 ^= (z4 << 15) & 0x9b868000u;
    return z4 >> (32 - bits);
  }
  double nextDouble() {  // BitsStreamGenerator.nextDouble
    const int64_t high = (int64_t)next(26) << 26;
    const int64_t low = (int64_t)next(26);
    return (double)(high | low) * 0x1.0p-52;
  }
};

// org.apache.commons.math3.distribution.PoissonDistribution.sample() for mean < 40.  Appendix A.4.
inline int poisson_sample(Well19937c& rng, double mean) {
  const double p = std::exp(-mean);
  long n = 0;
  double r = 1.0;
  while ((double)n < 1000 * mean) {
    r *= rng.nextDouble();
    if (r >= p) n++; else return (int)n;
  }
  return (int)n;
}

// org.apache.spark.util.random.SamplingUtils.reservoirSampleAndCount over 0 until F.  Appendix A.5.
void reservoir(int F, int k, int64_t seed, std::vector<int>& out) {
  out.resize(k);
  int i = 0;
  for (; i < k && i < F; i++) out[i] = i;
  if (i < k) { out.resize(i); return; }
  XORShiftRandom rand(seed);
  long l = i;
  for (int item = k; item < F; item++) {
    l++;
    const long replacement = (long)(rand.nextDouble() * (double)l);
    if (replacement < k) out[(int)replacement] = item;
  }
}

// Scala 2.12 immutable.HashMap iteration order over Int keys (more than 4 entries).  Appendix A.6.
inline uint32_t scala_improve(uint32_t hcode) {
  uint32_t h = hcode + ~(hcode << 9);
  h = h ^ (h >> 14);
  h = h + (h << 4);
  return h ^ (h >> 10);
}
inline uint64_t trie_key(uint32_t h) {  // lexicographic order of 5-bit chunks, LSB chunk first
  uint64_t k = 0;
  for (int lvl = 0; lvl < 7; lvl++) k = (k << 5) | ((h >> (5 * lvl)) & 31u);
  return k;
}
// Scala 2.12 mutable.HashMap (initial table 16) iteration order for <= 4 Int keys inserted in `ins` order.
inline uint32_t byteswap32(uint32_t v) {
  uint32_t hc = v * 0x9e3775cdu;
  hc = ((hc & 0xff) << 24) | ((hc & 0xff00) << 8) | ((hc >> 8) & 0xff00) | (hc >> 24);
  return hc * 0x9e3775cdu;
}
void scala_group_order(const std::vector<int>& ins, std::vector<int>& out) {
  out.clear();
  // step 1: iteration order of the mutable.HashMap the keys were inserted into
  std::vector<int> mut;
  {
    // table grows (x2) when size exceeds 75% load; order below assumes the final table size.
    int tbl = 16;
    while ((int)ins.size() > tbl * 3 / 4) tbl *= 2;
    const int ones = tbl - 1;
    const int seed = __builtin_popcount(ones);
    const int expo = __builtin_clz((unsigned)ones);
    std::vector<std::vector<int>> buckets(tbl);
    for (int key : ins) {
      uint32_t i = byteswap32((uint32_t)key);
      const int rot = seed % 32;
      uint32_t imp = rot ? ((i >> rot) | (i << (32 - rot))) : i;
      const int idx = (int)((imp >> expo) & (uint32_t)ones);
      buckets[idx].insert(buckets[idx].begin(), key);  // newest first
    }
    for (int b = tbl - 1; b >= 0; b--) for (int key : buckets[b]) mut.push_back(key);
  }
  if (mut.size() <= 4) { out = mut; return; }  // Map1..Map4 keep insertion order
  std::vector<std::pair<uint64_t, int>> ks;
  for (int key : mut) ks.push_back({trie_key(scala_improve((uint32_t)key)), key});
  std::sort(ks.begin(), ks.end());
  for (auto& p : ks) out.push_back(p.second);
}

// ----------------------------------------------------------------------------------------------
// Parameters / metadata  (DecisionTreeMetadata.buildMetadata — Appendix A.1)
// ----------------------------------------------------------------------------------------------
struct Params {
  int maxDepth = 5, maxBins = 32, minInstancesPerNode = 1, numTrees = 1;
  double minInfoGain = 0.0, minWeightFractionPerNode = 0.0, subsamplingRate = 1.0;
  int maxMemoryInMB = 256;
  int bootstrap = 0;
  int64_t seed = 0;
  int featuresPerNode = 0;  // resolved by the caller (featureSubsetStrategy table, Appendix A.1)
  int impurity = 0;         // 0 variance (regression), 1 gini, 2 entropy (binary classification: numClasses == 2)
};

struct Meta {
  int64_t N = 0;
  int F = 0;
  double weightedN = 0;
  int maxPossibleBins = 0;
  std::vector<int> arity;    // 0 => continuous
  std::vector<int> numBins;  // per feature
  int metaMaxBins = 0;       // numBins.max captured before findSplits shrinks continuous features
  int featuresPerNode = 0;
  bool subsampling = false;
  double minWeightPerNode = 0;
  int numSplits(int f) const { return numBins[f] - 1; }
  bool isCont(int f) const { return arity[f] == 0; }
};

int build_meta(int64_t N, int F, const int* arity, const Params& p, Meta& m) {
  m.N = N; m.F = F; m.weightedN = (double)N;
  m.maxPossibleBins = (int)std::min<int64_t>(p.maxBins, N);
  m.arity.assign(arity, arity + F);
  int maxCat = 0, maxCatF = -1;
  for (int f = 0; f < F; f++) if (arity[f] > maxCat) { maxCat = arity[f]; maxCatF = f; }
  if (maxCatF >= 0 && maxCat > m.maxPossibleBins) {
    char buf[512];
    snprintf(buf, sizeof buf,
             "requirement failed: DecisionTree requires maxBins (= %d) to be at least as large as the number of values in each "
             "categorical feature, but categorical feature %d has %d values. Consider removing this and other categorical "
             "features with a large number of values, or add more training examples.",
             m.maxPossibleBins, maxCatF, maxCat);
    g_err = buf;
    return -2;
  }
  m.numBins.assign(F, m.maxPossibleBins);
  for (int f = 0; f < F; f++) if (arity[f] > 1) m.numBins[f] = arity[f];
  m.metaMaxBins = *std::max_element(m.numBins.begin(), m.numBins.end());
  m.featuresPerNode = p.featuresPerNode > 0 ? p.featuresPerNode : F;
  m.subsampling = m.featuresPerNode != F;
  m.minWeightPerNode = p.minWeightFractionPerNode * m.weightedN;
  return 0;
}

// ----------------------------------------------------------------------------------------------
// A.2 findSplitsForContinuousFeature (RandomForest.scala).  `vals` = sampled values of one feature (all of them,
// zeros included); unit instance weights.
// ----------------------------------------------------------------------------------------------
void find_splits_feature(const double* vals, int64_t n, int64_t stride_elems, int numSplits, double weightedNumSamples,
                         std::vector<double>& out) {
  out.clear();
  std::vector<double> nz;
  nz.reserve(n);
  for (int64_t i = 0; i < n; i++) { double x = vals[i * stride_elems]; if (x != 0.0) nz.push_back(x); }
  if (nz.empty()) return;  // partValueCountMap.isEmpty
  const int64_t count = (int64_t)nz.size();
  std::sort(nz.begin(), nz.end());
  std::vector<std::pair<double, double>> vc;  // (value, weight-sum), ascending
  for (double x : nz) { if (!vc.empty() && vc.back().first == x) vc.back().second += 1.0; else vc.push_back({x, 1.0}); }
  double partNum = 0;
  for (auto& p : vc) partNum += p.second;
  const double tol = 0x1.0p-52 * (double)count * 100;
  if (weightedNumSamples - partNum > tol) {
    const double zc = weightedNumSamples - partNum;
    auto it = std::lower_bound(vc.begin(), vc.end(), std::make_pair(0.0, -1.0));
    vc.insert(it, {0.0, zc});
  }
  const int possible = (int)vc.size() - 1;
  if (possible == 0) return;
  if (possible <= numSplits) {
    for (int i = 1; i <= possible; i++) out.push_back((vc[i - 1].first + vc[i].first) / 2.0);
    return;
  }
  const double stride = weightedNumSamples / (double)(numSplits + 1);
  double cur = vc[0].second, target = stride;
  for (size_t i = 1; i < vc.size(); i++) {
    const double prev = cur;
    cur += vc[i].second;
    if (std::fabs(prev - target) < std::fabs(cur - target)) {
      // MLlib never emits more than numSplits here when the sample has its expected size; with this build's own
      // row sampler the sample can be larger than expected, so the count is capped (documented in DESIGN.md).
      if ((int)out.size() < numSplits) out.push_back((vc[i - 1].first + vc[i].first) / 2.0);
      target += stride;
    }
  }
}

// A.3 TreePoint.findBin
inline int find_bin(double x, const double* thr, int nthr) {
  // number of thresholds < x  (java.util.Arrays.binarySearch semantics: x == t_k -> k; NaN orders above every double -> last bin)
  if (x != x) return nthr;
  int lo = 0, hi = nthr;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (thr[mid] < x) lo = mid + 1; else hi = mid; }
  return lo;
}

// ----------------------------------------------------------------------------------------------
// Trees under construction (LearningNode) and finished (Node)
// ----------------------------------------------------------------------------------------------
// Regression (VarianceCalculator): stats = (weight, sum w*y, sum w*y*y) + rawCount.
// Classification (GiniCalculator / EntropyCalculator, numClasses == 2): stats = class counts; here cnt = count of class 0,
// sum = count of class 1, sq unused, raw = rawCount.  The weighted count of a calculator is stats.sum == cnt + sum.
struct Calc { double cnt = 0, sum = 0, sq = 0, raw = 0; };
static int g_impurity = 0;   // set by b2rf_oracle_fit for the duration of one fit (the oracle is test infrastructure, one fit at a time)
inline double calc_count(const Calc& c) { return g_impurity == 0 ? c.cnt : c.cnt + c.sum; }
inline double calc_impurity(const Calc& c) {
  if (g_impurity == 0) {               // Variance.calculate
    if (c.cnt == 0) return 0;
    const double sl = c.sq - (c.sum * c.sum) / c.cnt;
    return sl / c.cnt;
  }
  const double counts[2] = {c.cnt, c.sum};
  const double total = counts[0] + counts[1];
  if (total == 0) return 0;
  if (g_impurity == 1) {               // Gini.calculate(counts, totalCount)
    double imp = 1.0;
    for (int k = 0; k < 2; k++) { const double freq = counts[k] / total; imp -= freq * freq; }
    return imp;
  }
  double imp = 0.0;                    // Entropy.calculate: log2(x) = log(x) / log(2)
  for (int k = 0; k < 2; k++) if (counts[k] != 0) { const double freq = counts[k] / total; imp -= freq * (std::log(freq) / std::log(2.0)); }
  return imp;
}
// regression: mean label; classification: indexOfLargestArrayElement(stats) (first maximum), 0 for an empty calculator
inline double calc_predict(const Calc& c) {
  if (g_impurity == 0) return c.cnt == 0 ? 0.0 : c.sum / c.cnt;
  return c.sum > c.cnt ? 1.0 : 0.0;
}
// ordering key of a category in binsToBestSplit: regression -> predict; binary classification -> stats(1), the count of class 1
inline double calc_centroid(const Calc& c) { return g_impurity == 0 ? calc_predict(c) : c.sum; }

struct Stats {  // ImpurityStats
  double gain = 0, impurity = 0;
  Calc calc, left, right;
  bool valid = true, has = false;
};

struct LNode {
  int id = 1;  // heap index
  int left = -1, right = -1;
  bool isLeaf = false;
  bool hasSplit = false;
  int feature = -1;
  double threshold = 0;
  int splitBin = -1;             // continuous: bins <= splitBin go left
  std::vector<int> leftCats;     // categorical: bins in this list go left (sorted-by-centroid order)
  Stats stats;
};

struct FlatForest {
  std::vector<int64_t> tree_off;  // T+1
  std::vector<int> feature, left, right, ncat, cat_off, cat_len;
  std::vector<double> threshold, prediction, impurity, gain, stats;  // stats: 4 per node
  std::vector<double> cats;
};

struct TraceRec {  // one per node evaluated by binsToBestSplit, in processing order
  int group, tree, nodeId, feature, splitIdx;
  double gain, impurity, cnt, sum, sq;
};

struct Forest {
  Meta meta;
  Params params;
  std::vector<std::vector<double>> thresholds;
  std::vector<std::vector<LNode>> trees;
  FlatForest flat;
  std::vector<TraceRec> trace;
  std::vector<int> group_sizes;
  double t_init = 0, t_findSplits = 0, t_hist = 0, t_choose = 0, t_total = 0;
  int64_t updates = 0;
};

inline int level_of(int id) { return 31 - __builtin_clz((unsigned)id); }

// A.8: evaluate one candidate (calculateImpurityStats).  `st` is the running gainAndImpurityStats (has=false => null).
inline void calc_impurity_stats(Stats& st, const Calc& L, const Calc& R, const Meta& m, const Params& p) {
  Calc parent; double imp;
  if (!st.has) {
    parent.cnt = L.cnt + R.cnt; parent.sum = L.sum + R.sum; parent.sq = L.sq + R.sq; parent.raw = L.raw + R.raw;
    imp = calc_impurity(parent);
  } else { parent = st.calc; imp = st.impurity; }
  Stats out;
  out.has = true; out.calc = parent; out.impurity = imp;
  const double lc = calc_count(L), rcn = calc_count(R);
  const double total = lc + rcn;
  const bool bad = (L.raw < p.minInstancesPerNode) || (R.raw < p.minInstancesPerNode) || (lc < m.minWeightPerNode) ||
                   (rcn < m.minWeightPerNode);
  if (bad) { out.gain = -1.7976931348623157e308; out.valid = false; st = out; return; }
  const double li = calc_impurity(L), ri = calc_impurity(R);
  const double lw = lc / total, rw = rcn / total;
  const double gain = imp - lw * li - rw * ri;
  if (gain < p.minInfoGain) { out.gain = -1.7976931348623157e308; out.valid = false; st = out; return; }
  out.gain = gain; out.left = L; out.right = R; out.valid = true;
  st = out;
}

struct BestSplit {
  int feature = -1, splitIdx = -1;
  std::vector<int> leftCats;
  Stats stats;
};

// A.8 binsToBestSplit for one node.  hist: per subset position j, numBins[f] cells of 4 doubles.
void bins_to_best_split(double* hist, const std::vector<int64_t>& foff, const std::vector<int>& feats, const Meta& m,
                        const Params& p, const LNode& node, BestSplit& best) {
  Stats running;
  if (level_of(node.id) > 0) { running = node.stats; running.has = true; }
  bool any = false;
  double bestGain = 0;
  for (size_t j = 0; j < feats.size(); j++) {
    const int f = feats[j];
    const int ns = m.numSplits(f);
    if (ns == 0) continue;
    double* h = hist + foff[j] * 4;
    int fbest = -1; Stats fstats; double fgain = 0;
    std::vector<int> order;
    if (m.isCont(f)) {
      for (int s = 0; s < ns; s++) for (int k = 0; k < 4; k++) h[(s + 1) * 4 + k] += h[s * 4 + k];
      for (int s = 0; s < ns; s++) {
        Calc L{h[s * 4], h[s * 4 + 1], h[s * 4 + 2], h[s * 4 + 3]};
        Calc R{h[ns * 4] - L.cnt, h[ns * 4 + 1] - L.sum, h[ns * 4 + 2] - L.sq, h[ns * 4 + 3] - L.raw};
        calc_impurity_stats(running, L, R, m, p);
        if (fbest < 0 || running.gain > fgain) { fbest = s; fgain = running.gain; fstats = running; }
      }
    } else {
      const int nc = m.numBins[f];
      std::vector<std::pair<double, int>> cen(nc);
      for (int c = 0; c < nc; c++) {
        Calc cs{h[c * 4], h[c * 4 + 1], h[c * 4 + 2], h[c * 4 + 3]};
        cen[c] = {calc_count(cs) != 0 ? calc_centroid(cs) : 1.7976931348623157e308, c};
      }
      std::stable_sort(cen.begin(), cen.end(), [](const std::pair<double, int>& a, const std::pair<double, int>& b) { return a.first < b.first; });
      order.resize(nc);
      for (int c = 0; c < nc; c++) order[c] = cen[c].second;
      for (int s = 0; s < ns; s++) {
        const int curc = order[s], nxt = order[s + 1];
        for (int k = 0; k < 4; k++) h[nxt * 4 + k] += h[curc * 4 + k];
      }
      const int last = order[nc - 1];
      for (int s = 0; s < ns; s++) {
        const int fv = order[s];
        Calc L{h[fv * 4], h[fv * 4 + 1], h[fv * 4 + 2], h[fv * 4 + 3]};
        Calc R{h[last * 4] - L.cnt, h[last * 4 + 1] - L.sum, h[last * 4 + 2] - L.sq, h[last * 4 + 3] - L.raw};
        calc_impurity_stats(running, L, R, m, p);
        if (fbest < 0 || running.gain > fgain) { fbest = s; fgain = running.gain; fstats = running; }
      }
    }
    if (!any || fgain > bestGain) {
      any = true; bestGain = fgain;
      best.feature = f; best.splitIdx = fbest; best.stats = fstats;
      best.leftCats.clear();
      if (!m.isCont(f)) best.leftCats.assign(order.begin(), order.begin() + fbest + 1);
    }
  }
  if (!any) {  // no feature with splits: dummy split + invalid stats from parentStats (== node totals)
    best.feature = feats.empty() ? 0 : feats[0];
    best.splitIdx = 0;
    best.leftCats.clear();
    Stats s; s.has = true; s.valid = false; s.gain = -1.7976931348623157e308;
    if (running.has) s.calc = running.calc;
    else if (!feats.empty()) {  // parentStats == node totals == the single bin of any splitless feature
      const double* h = hist + foff[0] * 4;
      for (int b = 0; b < m.numBins[feats[0]]; b++) { s.calc.cnt += h[b * 4]; s.calc.sum += h[b * 4 + 1]; s.calc.sq += h[b * 4 + 2]; s.calc.raw += h[b * 4 + 3]; }
    }
    s.impurity = calc_impurity(s.calc);
    best.stats = s;
  }
}

inline bool goes_left_binned(const LNode& n, int bin, const Meta& m) {
  if (m.isCont(n.feature)) return bin <= n.splitBin;   // A.9 (bin != numSplits && thr[bin] <= threshold)
  for (int c : n.leftCats) if (c == bin) return true;
  return false;
}

double now_s() {
#ifdef _OPENMP
  return omp_get_wtime();
#else
  return (double)clock() / CLOCKS_PER_SEC;
#endif
}

// toNode(prune=true) -> flat pre-order arrays.  Appendix A.10.
struct FNode { int feature, left, right, ncat; double thr, pred, imp, gain; Calc calc; std::vector<int> cats; };
int flatten(const std::vector<LNode>& t, int i, const Meta& m, std::vector<FNode>& out) {
  const LNode& n = t[i];
  const int me = (int)out.size();
  out.push_back(FNode());
  if (n.left >= 0) {
    const int l = flatten(t, n.left, m, out);
    const int r = flatten(t, n.right, m, out);
    const bool lleaf = out[l].feature < 0, rleaf = out[r].feature < 0;
    if (lleaf && rleaf && out[l].pred == out[r].pred) {
      const double lp = out[l].pred;
      out.resize(me + 1);
      FNode& o = out[me];
      o.feature = -1; o.left = o.right = -1; o.ncat = 0; o.thr = 0; o.pred = lp;
      o.imp = n.stats.impurity; o.gain = 0; o.calc = n.stats.calc;
      return me;
    }
    FNode& o = out[me];
    o.feature = n.feature; o.left = l; o.right = r; o.thr = n.threshold;
    o.ncat = m.isCont(n.feature) ? 0 : m.numBins[n.feature];
    o.cats = n.leftCats;
    o.pred = calc_predict(n.stats.calc); o.imp = n.stats.impurity; o.gain = n.stats.gain; o.calc = n.stats.calc;
    return me;
  }
  FNode& o = out[me];
  o.feature = -1; o.left = o.right = -1; o.ncat = 0; o.thr = 0;
  o.pred = calc_predict(n.stats.calc);
  o.imp = n.stats.valid ? n.stats.impurity : -1.0;
  o.gain = 0; o.calc = n.stats.calc;
  return me;
}

void export_flat(Forest& fo) {
  FlatForest& ff = fo.flat;
  ff = FlatForest();
  ff.tree_off.push_back(0);
  for (auto& t : fo.trees) {
    std::vector<FNode> nodes;
    flatten(t, 0, fo.meta, nodes);  // node indices are relative to the tree (pre-order)
    for (auto& n : nodes) {
      ff.feature.push_back(n.feature); ff.left.push_back(n.left); ff.right.push_back(n.right);
      ff.threshold.push_back(n.thr); ff.prediction.push_back(n.pred); ff.impurity.push_back(n.imp); ff.gain.push_back(n.gain);
      ff.stats.push_back(n.calc.cnt); ff.stats.push_back(n.calc.sum); ff.stats.push_back(n.calc.sq); ff.stats.push_back(n.calc.raw);
      ff.ncat.push_back(n.ncat); ff.cat_off.push_back((int)ff.cats.size()); ff.cat_len.push_back((int)n.cats.size());
      for (int c : n.cats) ff.cats.push_back((double)c);
    }
    ff.tree_off.push_back((int64_t)ff.feature.size());
  }
}

}  // namespace

// ================================================================================================
// C API (host pointers only)
// ================================================================================================
extern "C" {

const char* b2rf_oracle_last_error() { return g_err.c_str(); }

// ---- RNG known-answer hooks (Appendix B) ----
void b2rf_oracle_java_random(int64_t seed, int32_t* next_int, int64_t* next_long) {
  JavaRandom a(seed); *next_int = a.nextInt();
  JavaRandom b(seed); *next_long = b.nextLong();
}
void b2rf_oracle_java_random_longs(int64_t seed, int n, int64_t* out) { JavaRandom r(seed); for (int i = 0; i < n; i++) out[i] = r.nextLong(); }
void b2rf_oracle_xorshift_doubles(int64_t seed, int n, double* out) { XORShiftRandom r(seed); for (int i = 0; i < n; i++) out[i] = r.nextDouble(); }
int32_t b2rf_oracle_xorshift_nextint(int64_t seed) { XORShiftRandom r(seed); return r.nextInt(); }
void b2rf_oracle_well_doubles(int64_t seed, int n, double* out) { Well19937c r(seed); for (int i = 0; i < n; i++) out[i] = r.nextDouble(); }
void b2rf_oracle_poisson(int64_t seed, double mean, int n, int32_t* out) { Well19937c r(seed); for (int i = 0; i < n; i++) out[i] = poisson_sample(r, mean); }
void b2rf_oracle_reservoir(int F, int k, int64_t seed, int32_t* out) { std::vector<int> o; reservoir(F, k, seed, o); for (size_t i = 0; i < o.size(); i++) out[i] = o[i]; }
void b2rf_oracle_group_order(const int32_t* keys, int n, int32_t* out) {
  std::vector<int> in(keys, keys + n), o; scala_group_order(in, o); for (int i = 0; i < n; i++) out[i] = o[i];
}

// ---- A.2 ---- sampleX: [n_s][F] row-major. thresholds_out: [F][maxSplits]; nsplits_out[F].
// numSplitsWanted[f] = metadata.numSplits(f) before findSplits (maxPossibleBins-1 for continuous).
int b2rf_oracle_find_splits(const double* sampleX, int64_t n_s, int F, const int32_t* arity, int maxPossibleBins,
                            double weightedNumSamples, double* thresholds_out, int32_t* nsplits_out) {
  const int maxSplits = maxPossibleBins - 1;
  std::vector<double> thr;
  for (int f = 0; f < F; f++) {
    if (arity[f] != 0) { nsplits_out[f] = 0; continue; }
    find_splits_feature(sampleX + f, n_s, F, maxSplits, weightedNumSamples, thr);
    nsplits_out[f] = (int)thr.size();
    for (size_t i = 0; i < thr.size(); i++) thresholds_out[(int64_t)f * maxSplits + i] = thr[i];
  }
  return 0;
}

// The findSplits row sample used when fraction < 1.  MLlib uses RDD.sample (partition-wise Bernoulli), which is
// not reproducible outside Spark; this build's contract replaces it by a per-global-row hash that oracle
// and GPU share.  sampleSeed = new XORShiftRandom(seed).nextInt() as in RandomForest.findSplits.
static inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
int b2rf_oracle_sample_keep(int64_t seed, int64_t global_row, double fraction) {
  if (fraction >= 1.0) return 1;
  const int32_t ss = XORShiftRandom(seed).nextInt();
  const uint64_t h = mix64(((uint64_t)(uint32_t)ss << 32 | 0x9E3779B9u) + (uint64_t)global_row * 0x9E3779B97F4A7C15ULL);
  return ((double)(h >> 11) * 0x1.0p-53) < fraction;
}

// ---- A.3 ---- bins_out int32 [N][F] row-major (MLlib TreePoint layout). returns -3 on invalid categorical value.
int b2rf_oracle_bin(const double* X, int64_t N, int F, const double* thresholds, const int32_t* nsplits, const int32_t* arity,
                    int maxSplits, int32_t* bins_out) {
  for (int64_t i = 0; i < N; i++)
    for (int f = 0; f < F; f++) {
      const double x = X[i * F + f];
      if (arity[f] == 0) bins_out[i * F + f] = find_bin(x, thresholds + (int64_t)f * maxSplits, nsplits[f]);
      else {
        if (x < 0 || x >= arity[f]) {
          char buf[256];
          snprintf(buf, sizeof buf, "DecisionTree given invalid data: Feature %d is categorical with values in {0,...,%d}, but a data point gives it value %g.", f, arity[f] - 1, x);
          g_err = buf; return -3;
        }
        bins_out[i * F + f] = (int)x;
      }
    }
  return 0;
}

// ---- A.4 ---- counts_out uint8 [N][T] row-major (row-major draw order).
int b2rf_oracle_bag(int64_t seed, const int64_t* part_offsets, int P, int T, double rate, int bootstrap, uint8_t* counts_out) {
  for (int p = 0; p < P; p++) {
    const int64_t r0 = part_offsets[p], r1 = part_offsets[p + 1];
    if (bootstrap) {
      Well19937c rng(seed + p + 1);
      for (int64_t i = r0; i < r1; i++) for (int t = 0; t < T; t++) { int c = poisson_sample(rng, rate); counts_out[i * T + t] = (uint8_t)std::min(c, 255); }
    } else if (T == 1 && rate == 1.0) {
      for (int64_t i = r0; i < r1; i++) counts_out[i * T] = 1;
    } else {
      XORShiftRandom rng(seed + p + 1);
      for (int64_t i = r0; i < r1; i++) for (int t = 0; t < T; t++) counts_out[i * T + t] = rng.nextDouble() < rate ? 1 : 0;
    }
  }
  return 0;
}

// ---- A.7 stand-alone: histogram of one node group given precomputed slots.  bins int32 [N][F];
// slot int32 [T_g][N] (-1 = not in group); counts uint8 [N][T] ; tree_ids[T_g] ; feats int32 [slots][Fsub];
// cell offsets: off[slots][Fsub] (in cells); hist_out double [cells][4].
int b2rf_oracle_hist(const int32_t* bins, int64_t N, int F, const double* y, const uint8_t* counts, int T, const int32_t* tree_ids, int Tg,
                     const int32_t* slot, const int32_t* feats, int Fsub, const int64_t* off, double* hist_out) {
  for (int64_t i = 0; i < N; i++)
    for (int g = 0; g < Tg; g++) {
      const int s = slot[(int64_t)g * N + i];
      if (s < 0) continue;
      const double w = (double)counts[i * T + tree_ids[g]];
      const double yy = y[i];
      for (int j = 0; j < Fsub; j++) {
        const int f = feats[(int64_t)s * Fsub + j];
        double* c = hist_out + (off[(int64_t)s * Fsub + j] + bins[i * F + f]) * 4;
        c[0] += w; c[1] += w * yy; c[2] += w * yy * yy; c[3] += w;
      }
    }
  return 0;
}

// ---- full fit ----
struct B2rfOracleParams {
  int32_t maxDepth, maxBins, minInstancesPerNode, numTrees, featuresPerNode, maxMemoryInMB, bootstrap, numThreads;
  double minInfoGain, minWeightFractionPerNode, subsamplingRate;
  int64_t seed;
  int32_t impurity;   // 0 variance, 1 gini, 2 entropy (labels must be 0.0 / 1.0)
  int32_t reserved;
};

// X [N][F] row-major f64; arity[F] (0 = continuous); part_offsets[P+1]; sample_idx: rows used by findSplits (n_sample of them,
// ascending) or NULL => all rows (fraction 1).  fraction: samplesFractionForFindSplits value to use.
int b2rf_oracle_fit(const double* X, int64_t N, int F, const double* y, const int32_t* arity, const B2rfOracleParams* pp,
                    const int64_t* part_offsets, int P, const int64_t* sample_idx, int64_t n_sample, double fraction, void** out) {
  g_err.clear();
  Forest* fo = new Forest();
  Params& p = fo->params;
  p.maxDepth = pp->maxDepth; p.maxBins = pp->maxBins; p.minInstancesPerNode = pp->minInstancesPerNode; p.numTrees = pp->numTrees;
  p.featuresPerNode = pp->featuresPerNode; p.maxMemoryInMB = pp->maxMemoryInMB; p.bootstrap = pp->bootstrap;
  p.minInfoGain = pp->minInfoGain; p.minWeightFractionPerNode = pp->minWeightFractionPerNode; p.subsamplingRate = pp->subsamplingRate;
  p.seed = pp->seed; p.impurity = pp->impurity;
  if (p.impurity < 0 || p.impurity > 2) { g_err = "impurity must be 0 (variance), 1 (gini) or 2 (entropy)"; delete fo; return -1; }
  if (p.impurity != 0) for (int64_t i = 0; i < N; i++) if (!(y[i] == 0.0 || y[i] == 1.0)) { g_err = "classification labels must be 0.0 or 1.0 (numClasses == 2)"; delete fo; return -1; }
  g_impurity = p.impurity;
  const int nthreads = std::max(1, pp->numThreads);
  const double t0 = now_s();
  Meta& m = fo->meta;
  int rc = build_meta(N, F, arity, p, m);
  if (rc) { delete fo; return rc; }
  const int T = p.numTrees;
  const int maxSplits = m.maxPossibleBins - 1;

  // --- findSplits ---
  const double t1 = now_s();
  fo->thresholds.assign(F, {});
  {
    std::vector<double> samp;
    const double* sx = X; int64_t ns = N;
    if (sample_idx) {
      samp.resize((size_t)n_sample * F);
      for (int64_t i = 0; i < n_sample; i++) memcpy(&samp[(size_t)i * F], X + sample_idx[i] * F, sizeof(double) * F);
      sx = samp.data(); ns = n_sample;
    }
    const double wns = fraction * m.weightedN;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic)
    for (int f = 0; f < F; f++) {
      if (!m.isCont(f)) continue;
      find_splits_feature(sx + f, ns, F, maxSplits, wns, fo->thresholds[f]);
    }
    for (int f = 0; f < F; f++) if (m.isCont(f)) m.numBins[f] = (int)fo->thresholds[f].size() + 1;
  }
  const double t2 = now_s();

  // --- TreePoint + BaggedPoint ---
  std::vector<int32_t> bins((size_t)N * F);
  {
    int err = 0;
#pragma omp parallel for num_threads(nthreads)
    for (int64_t i = 0; i < N; i++)
      for (int f = 0; f < F; f++) {
        const double x = X[i * F + f];
        if (m.isCont(f)) bins[i * F + f] = find_bin(x, fo->thresholds[f].data(), (int)fo->thresholds[f].size());
        else { if (x < 0 || x >= m.arity[f]) err = f + 1; bins[i * F + f] = (int)x; }
      }
    if (err) {
      char buf[256]; snprintf(buf, sizeof buf, "DecisionTree given invalid data: Feature %d is categorical with values in {0,...,%d}, but a data point gives it an out-of-range value.", err - 1, m.arity[err - 1] - 1);
      g_err = buf; delete fo; return -3;
    }
  }
  std::vector<uint8_t> counts((size_t)N * T);
  {
#pragma omp parallel for num_threads(nthreads) schedule(dynamic)
    for (int pi = 0; pi < P; pi++) {
      int64_t po[2] = {part_offsets[pi], part_offsets[pi + 1]};
      // seed offset: partition index pi -> seed + pi + 1
      const int64_t r0 = po[0], r1 = po[1];
      if (p.bootstrap) {
        Well19937c rng(p.seed + pi + 1);
        for (int64_t i = r0; i < r1; i++) for (int t = 0; t < T; t++) counts[i * T + t] = (uint8_t)std::min(poisson_sample(rng, p.subsamplingRate), 255);
      } else if (T == 1 && p.subsamplingRate == 1.0) {
        for (int64_t i = r0; i < r1; i++) counts[i * T] = 1;
      } else {
        XORShiftRandom rng(p.seed + pi + 1);
        for (int64_t i = r0; i < r1; i++) for (int t = 0; t < T; t++) counts[i * T + t] = rng.nextDouble() < p.subsamplingRate ? 1 : 0;
      }
    }
  }
  const double t3 = now_s();
  fo->t_init = (t1 - t0) + (t3 - t2);
  fo->t_findSplits = t2 - t1;

  // --- RandomForest.run main loop (A.6) ---
  fo->trees.assign(T, {});
  for (int t = 0; t < T; t++) { fo->trees[t].reserve(64); fo->trees[t].push_back(LNode()); }
  std::vector<std::pair<int, int>> stack;  // (tree, node index in trees[t]); back() is the head
  for (int t = 0; t < T; t++) stack.push_back({t, 0});
  JavaRandom rng(p.seed);
  const int64_t maxMem = (int64_t)p.maxMemoryInMB * 1024 * 1024;
  std::vector<int> allFeats(F);
  for (int f = 0; f < F; f++) allFeats[f] = f;
  int64_t allBins = 0;
  for (int f = 0; f < F; f++) allBins += m.numBins[f];

  struct GNode { int tree, nidx; std::vector<int> feats; std::vector<int64_t> foff; int64_t base; int64_t cells; };
  int groupNo = 0;
  std::vector<std::vector<double>> tbuf(nthreads);
  while (!stack.empty()) {
    // selectNodesToSplit
    std::vector<GNode> group;
    std::vector<int> treeInsertOrder;
    int64_t mem = 0;
    while (!stack.empty()) {
      auto [t, ni] = stack.back();
      GNode g; g.tree = t; g.nidx = ni;
      int64_t totalBins;
      if (m.subsampling) {
        reservoir(F, m.featuresPerNode, rng.nextLong(), g.feats);
        totalBins = 0; for (int f : g.feats) totalBins += m.numBins[f];
      } else { g.feats = allFeats; totalBins = allBins; }
      const int64_t nodeMem = (p.impurity == 0 ? 3 : 2) * totalBins * 8;   // aggregateSizeForNode: 3 (regression) or numClasses (= 2)
      if (mem + nodeMem <= maxMem || mem == 0) {
        stack.pop_back();
        if (std::find(treeInsertOrder.begin(), treeInsertOrder.end(), t) == treeInsertOrder.end()) treeInsertOrder.push_back(t);
        g.cells = totalBins;
        group.push_back(std::move(g));
        mem += nodeMem;
      } else break;
    }
    int64_t totalCells = 0;
    for (auto& g : group) {
      g.base = totalCells;
      g.foff.resize(g.feats.size());
      int64_t o = 0;
      for (size_t j = 0; j < g.feats.size(); j++) { g.foff[j] = o; o += m.numBins[g.feats[j]]; }
      totalCells += g.cells;
    }
    fo->group_sizes.push_back((int)group.size());
    // per tree: heap id -> group slot
    std::vector<std::map<int, int>> slotOf(T);
    std::vector<char> treeIn(T, 0);
    for (size_t s = 0; s < group.size(); s++) { slotOf[group[s].tree][fo->trees[group[s].tree][group[s].nidx].id] = (int)s; treeIn[group[s].tree] = 1; }
    std::vector<int> gtrees;
    for (int t = 0; t < T; t++) if (treeIn[t]) gtrees.push_back(t);

    // --- aggregation (A.7): per-thread partial aggregators over contiguous partition ranges, merged in order ---
    const double th0 = now_s();
    for (int k = 0; k < nthreads; k++) tbuf[k].assign((size_t)totalCells * 4, 0.0);
    int64_t upd = 0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : upd)
    for (int k = 0; k < nthreads; k++) {
      double* H = tbuf[k].data();
      const int p0 = (int)((int64_t)P * k / nthreads), p1 = (int)((int64_t)P * (k + 1) / nthreads);
      for (int64_t i = part_offsets[p0]; i < part_offsets[p1]; i++) {
        const int32_t* b = &bins[i * F];
        const double yy = y[i];
        for (int t : gtrees) {
          // predictImpl from the root
          const std::vector<LNode>& tr = fo->trees[t];
          int ni = 0; int id;
          for (;;) {
            const LNode& n = tr[ni];
            if (n.isLeaf || !n.hasSplit) { id = n.id; break; }
            const bool l = goes_left_binned(n, b[n.feature], m);
            if (n.left < 0) { id = l ? 2 * n.id : 2 * n.id + 1; break; }
            ni = l ? n.left : n.right;
          }
          auto it = slotOf[t].find(id);
          if (it == slotOf[t].end()) continue;
          const GNode& g = group[it->second];
          const double w = (double)counts[i * T + t];
          const double wy = w * yy, wyy = w * yy * yy;
          const int cls = (int)yy;   // classification: allStats(offset + label) += numSamples * sampleWeight
          double* base = H + g.base * 4;
          const int nf = (int)g.feats.size();
          for (int j = 0; j < nf; j++) {
            double* c = base + (g.foff[j] + b[g.feats[j]]) * 4;
            if (p.impurity == 0) { c[0] += w; c[1] += wy; c[2] += wyy; c[3] += w; }
            else { c[cls] += w; c[3] += w; }
          }
          upd += nf;
        }
      }
    }
    for (int k = 1; k < nthreads; k++) {
      double* a = tbuf[0].data(); const double* bsrc = tbuf[k].data();
      const int64_t n4 = totalCells * 4;
#pragma omp parallel for num_threads(nthreads)
      for (int64_t q = 0; q < n4; q++) a[q] += bsrc[q];
    }
    fo->updates += upd;
    const double th1 = now_s();
    fo->t_hist += th1 - th0;

    // --- binsToBestSplit per node (A.8) ---
    std::vector<BestSplit> bests(group.size());
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 16)
    for (int64_t s = 0; s < (int64_t)group.size(); s++) {
      const GNode& g = group[s];
      bins_to_best_split(tbuf[0].data() + g.base * 4, g.foff, g.feats, m, p, fo->trees[g.tree][g.nidx], bests[s]);
    }
    // --- driver tail (A.6/A.10): iterate trees in Scala map order, nodes in pop order ---
    std::vector<int> order;
    scala_group_order(treeInsertOrder, order);
    for (int t : order) {
      for (size_t s = 0; s < group.size(); s++) {
        if (group[s].tree != t) continue;
        const BestSplit& bs = bests[s];
        std::vector<LNode>& tr = fo->trees[t];
        const int ni = group[s].nidx;
        const int id = tr[ni].id;
        const int lvl = level_of(id);
        const bool isLeaf = (bs.stats.gain <= 0) || (lvl == p.maxDepth);
        tr[ni].isLeaf = isLeaf;
        tr[ni].stats = bs.stats;
        fo->trace.push_back({groupNo, t, id, bs.feature, bs.splitIdx, bs.stats.gain, bs.stats.impurity, bs.stats.calc.cnt, bs.stats.calc.sum, bs.stats.calc.sq});
        if (!isLeaf) {
          tr[ni].hasSplit = true;
          tr[ni].feature = bs.feature;
          if (m.isCont(bs.feature)) { tr[ni].splitBin = bs.splitIdx; tr[ni].threshold = fo->thresholds[bs.feature][bs.splitIdx]; }
          else tr[ni].leftCats = bs.leftCats;
          const bool childLeaf = (lvl + 1) == p.maxDepth;
          const double li = calc_impurity(bs.stats.left), ri = calc_impurity(bs.stats.right);
          const bool lLeaf = childLeaf || (std::fabs(li) < 0x1.0p-52);
          const bool rLeaf = childLeaf || (std::fabs(ri) < 0x1.0p-52);
          LNode L, Rn;
          L.id = 2 * id; L.isLeaf = lLeaf; L.stats.has = true; L.stats.valid = true; L.stats.gain = NAN; L.stats.impurity = li; L.stats.calc = bs.stats.left;
          Rn.id = 2 * id + 1; Rn.isLeaf = rLeaf; Rn.stats.has = true; Rn.stats.valid = true; Rn.stats.gain = NAN; Rn.stats.impurity = ri; Rn.stats.calc = bs.stats.right;
          const int li_idx = (int)tr.size(); tr.push_back(L);
          const int ri_idx = (int)tr.size(); tr.push_back(Rn);
          tr[ni].left = li_idx; tr[ni].right = ri_idx;
          if (!lLeaf) stack.push_back({t, li_idx});
          if (!rLeaf) stack.push_back({t, ri_idx});
        }
      }
    }
    fo->t_choose += now_s() - th1;
    groupNo++;
  }
  export_flat(*fo);
  fo->t_total = now_s() - t0;
  *out = fo;
  return 0;
}

void b2rf_oracle_free(void* h) { delete (Forest*)h; }

int64_t b2rf_oracle_num_nodes(void* h) { return (int64_t)((Forest*)h)->flat.feature.size(); }
int64_t b2rf_oracle_num_cats(void* h) { return (int64_t)((Forest*)h)->flat.cats.size(); }
int32_t b2rf_oracle_num_groups(void* h) { return (int32_t)((Forest*)h)->group_sizes.size(); }
int64_t b2rf_oracle_num_trace(void* h) { return (int64_t)((Forest*)h)->trace.size(); }
int64_t b2rf_oracle_updates(void* h) { return ((Forest*)h)->updates; }
void b2rf_oracle_timers(void* h, double* out5) {
  Forest* f = (Forest*)h; out5[0] = f->t_init; out5[1] = f->t_findSplits; out5[2] = f->t_hist; out5[3] = f->t_choose; out5[4] = f->t_total;
}
void b2rf_oracle_export(void* h, int64_t* tree_off, int32_t* feature, double* threshold, int32_t* left, int32_t* right, double* prediction,
                        double* impurity, double* gain, double* stats4, int32_t* ncat, int32_t* cat_off, int32_t* cat_len, double* cats) {
  const FlatForest& f = ((Forest*)h)->flat;
  const size_t n = f.feature.size();
  memcpy(tree_off, f.tree_off.data(), f.tree_off.size() * 8);
  memcpy(feature, f.feature.data(), n * 4); memcpy(threshold, f.threshold.data(), n * 8);
  memcpy(left, f.left.data(), n * 4); memcpy(right, f.right.data(), n * 4);
  memcpy(prediction, f.prediction.data(), n * 8); memcpy(impurity, f.impurity.data(), n * 8); memcpy(gain, f.gain.data(), n * 8);
  memcpy(stats4, f.stats.data(), n * 32);
  memcpy(ncat, f.ncat.data(), n * 4); memcpy(cat_off, f.cat_off.data(), n * 4); memcpy(cat_len, f.cat_len.data(), n * 4);
  if (!f.cats.empty()) memcpy(cats, f.cats.data(), f.cats.size() * 8);
}
// thresholds_out [F][maxPossibleBins-1], nsplits_out[F]
void b2rf_oracle_export_splits(void* h, double* thresholds_out, int32_t* nsplits_out) {
  Forest* f = (Forest*)h; const int ms = f->meta.maxPossibleBins - 1;
  for (int j = 0; j < f->meta.F; j++) {
    nsplits_out[j] = (int)f->thresholds[j].size();
    for (size_t i = 0; i < f->thresholds[j].size(); i++) thresholds_out[(int64_t)j * ms + i] = f->thresholds[j][i];
  }
}
void b2rf_oracle_export_groups(void* h, int32_t* sizes) { Forest* f = (Forest*)h; memcpy(sizes, f->group_sizes.data(), f->group_sizes.size() * 4); }
// trace: ints [n][5] = group, tree, nodeId, feature, splitIdx ; doubles [n][5] = gain, impurity, cnt, sum, sq
void b2rf_oracle_export_trace(void* h, int32_t* ints, double* dbls) {
  Forest* f = (Forest*)h;
  for (size_t i = 0; i < f->trace.size(); i++) {
    const TraceRec& r = f->trace[i];
    ints[i * 5] = r.group; ints[i * 5 + 1] = r.tree; ints[i * 5 + 2] = r.nodeId; ints[i * 5 + 3] = r.feature; ints[i * 5 + 4] = r.splitIdx;
    dbls[i * 5] = r.gain; dbls[i * 5 + 1] = r.impurity; dbls[i * 5 + 2] = r.cnt; dbls[i * 5 + 3] = r.sum; dbls[i * 5 + 4] = r.sq;
  }
}

// ---- A.9/A.10 predict on raw features: forest mean in tree order ----
int b2rf_oracle_predict(void* h, const double* X, int64_t N, int F, double* out) {
  const FlatForest& f = ((Forest*)h)->flat;
  const int T = (int)f.tree_off.size() - 1;
  for (int64_t i = 0; i < N; i++) {
    double acc = 0;
    for (int t = 0; t < T; t++) {
      int64_t base = f.tree_off[t]; int n = 0;
      while (f.feature[base + n] >= 0) {
        const int ft = f.feature[base + n];
        const double x = X[i * F + ft];
        bool l;
        if (f.ncat[base + n] == 0) l = x <= f.threshold[base + n];
        else { l = false; for (int c = 0; c < f.cat_len[base + n]; c++) if (f.cats[f.cat_off[base + n] + c] == x) { l = true; break; } }
        n = l ? f.left[base + n] : f.right[base + n];
      }
      acc += f.prediction[base + n];
    }
    out[i] = acc / T;
  }
  return 0;
}

// ---- classification: rawPrediction.  mode 0: RandomForestClassificationModel.predictRaw (votes(k) += classCounts(k) / total per
// tree, in tree order); mode 1: DecisionTreeClassificationModel.predictRaw (the leaf's class counts of tree 0) ----
int b2rf_oracle_predict_raw(void* h, const double* X, int64_t N, int F, int mode, double* out2) {
  const FlatForest& f = ((Forest*)h)->flat;
  const int T = (int)f.tree_off.size() - 1;
  for (int64_t i = 0; i < N; i++) {
    double v0 = 0, v1 = 0;
    for (int t = 0; t < (mode == 1 ? 1 : T); t++) {
      int64_t base = f.tree_off[t]; int n = 0;
      while (f.feature[base + n] >= 0) {
        const int ft = f.feature[base + n];
        const double x = X[i * F + ft];
        bool l;
        if (f.ncat[base + n] == 0) l = x <= f.threshold[base + n];
        else { l = false; for (int c = 0; c < f.cat_len[base + n]; c++) if (f.cats[f.cat_off[base + n] + c] == x) { l = true; break; } }
        n = l ? f.left[base + n] : f.right[base + n];
      }
      const double c0 = f.stats[(base + n) * 4], c1 = f.stats[(base + n) * 4 + 1];
      if (mode == 1) { v0 = c0; v1 = c1; }
      else { const double total = c0 + c1; if (total != 0) { v0 += c0 / total; v1 += c1 / total; } }
    }
    out2[i * 2] = v0; out2[i * 2 + 1] = v1;
  }
  return 0;
}

}  // extern "C"
