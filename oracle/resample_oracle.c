/* CPU oracle (plain C) for the integer/index part of the resampling path -- TEST INFRASTRUCTURE ONLY.
 *
 * Restates, without sharing code with it, the contract of the reference's legacy extension
 *   torchsearchsorted/src/cpu/searchsorted_cpu_wrapper.cpp:4-122  (row-batched binary search, both sides,
 *   1-row broadcast of either operand, int64 result)
 * which its only automated test pins to numpy.searchsorted (torchsearchsorted/test/test_searchsorted.py:34-44),
 * and the index step of sample_pdf (models/rendering.py:54-56: searchsorted(cdf, u, right=True), below/above clamp).
 * Pinned by tests/test_oracle_golden.py against tests/golden/g8_searchsorted.npz (numpy.searchsorted, which is
 * the reference's own test oracle) and g7_sample_pdf.npz (indices produced by the imported reference ops).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may load this.
 */
#include <stdint.h>

/* out[r*V + c] = number of elements of a[row r] that are < v (side_left) or <= v (right) */
void oracle_searchsorted(const float* a, int64_t rows_a, int64_t A, const float* v, int64_t rows_v, int64_t V,
                         int side_left, int64_t* out) {
  const int64_t rows = rows_a > rows_v ? rows_a : rows_v;
  for (int64_t r = 0; r < rows; ++r) {
    const float* ar = a + (rows_a == 1 ? 0 : r) * A;
    const float* vr = v + (rows_v == 1 ? 0 : r) * V;
    for (int64_t c = 0; c < V; ++c) {
      const float x = vr[c];
      int64_t n = 0;                       /* linear count: deliberately not a binary search */
      if (side_left) { while (n < A && ar[n] < x) ++n; }
      else           { while (n < A && ar[n] <= x) ++n; }
      out[r * V + c] = n;
    }
  }
}

/* inds, below, above of models/rendering.py:54-56 for one (cdf, u) pair set */
void oracle_pdf_indices(const float* cdf, const float* u, int64_t n_rays, int64_t n_cdf, int64_t n_u,
                        int64_t* inds, int64_t* below, int64_t* above) {
  oracle_searchsorted(cdf, n_rays, n_cdf, u, n_rays, n_u, 0, inds);
  for (int64_t i = 0; i < n_rays * n_u; ++i) {
    below[i] = inds[i] - 1 < 0 ? 0 : inds[i] - 1;
    above[i] = inds[i] > n_cdf - 1 ? n_cdf - 1 : inds[i];
  }
}
