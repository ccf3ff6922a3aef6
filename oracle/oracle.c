/*
 * oracle.c — CPU restatement of the reference's kNN search path.  TEST INFRASTRUCTURE ONLY
 * (see oracle.h for who may use it and for the parity-pinning status of each part).
 *
 * What it follows in /root/reference:
 *   src/lance_index.cpp:442-465            guards, label -> row id gather (done in the host mirror)
 *   rust_lib/src/lance_manager.rs:394-451  search(k, nprobes, refine_factor): ascending _distance, f32
 *   rust_lib/src/lance_manager.rs:461-471  deletes are filters applied before top-k
 *   rust_lib/src/lance_manager.rs:483-515  IVF_PQ parameters, metric mapping
 *   duckdb/extension/core_functions/include/core_functions/array_kernels.hpp:11-105
 *                                          sequential fp32 folds (the path that runs today, F4)
 * and, for what is not on the box (lance-index / lance-linalg 0.22.0), the published algorithm:
 * IVF probe by centroid distance, per-partition residual PQ lookup tables, ADC scan, refine.
 */
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int oracle_metric_from_string(const char *m) {
	if (!m) return ORACLE_L2;
	if (!strcmp(m, "cosine")) return ORACLE_COSINE;
	if (!strcmp(m, "dot") || !strcmp(m, "ip")) return ORACLE_DOT;
	return ORACLE_L2;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}

/* ------------------------------------------------------------------ distances */

/* no FMA contraction here: DuckDB's fold is a separate multiply and add (array_kernels.hpp:22-26),
 * and oracle/_ref checks this function bit for bit against it */
__attribute__((optimize("fp-contract=off"))) float oracle_distance_f32(int metric, const float *q,
                                                                      const float *x, int d) {
	if (metric == ORACLE_L2) {
		float acc = 0.f;
		for (int i = 0; i < d; i++) {
			float diff = q[i] - x[i];
			acc += diff * diff;
		}
		return acc;
	}
	if (metric == ORACLE_DOT) {
		float acc = 0.f;
		for (int i = 0; i < d; i++) acc += q[i] * x[i];
		return 1.0f - acc;
	}
	float dot = 0.f, nq = 0.f, nx = 0.f;
	for (int i = 0; i < d; i++) {
		dot += q[i] * x[i];
		nq += q[i] * q[i];
		nx += x[i] * x[i];
	}
	float den = sqrtf(nq * nx);
	if (!(den > 0.f)) return 1.0f;
	float sim = dot / den;
	if (sim > 1.f) sim = 1.f;
	if (sim < -1.f) sim = -1.f;
	return 1.0f - sim;
}

double oracle_distance_f64(int metric, const float *q, const float *x, int d) {
	if (metric == ORACLE_L2) {
		double acc = 0.0;
		for (int i = 0; i < d; i++) {
			double diff = (double)q[i] - (double)x[i];
			acc += diff * diff;
		}
		return acc;
	}
	if (metric == ORACLE_DOT) {
		double acc = 0.0;
		for (int i = 0; i < d; i++) acc += (double)q[i] * (double)x[i];
		return 1.0 - acc;
	}
	double dot = 0.0, nq = 0.0, nx = 0.0;
	for (int i = 0; i < d; i++) {
		dot += (double)q[i] * (double)x[i];
		nq += (double)q[i] * (double)q[i];
		nx += (double)x[i] * (double)x[i];
	}
	double den = sqrt(nq) * sqrt(nx);
	if (!(den > 0.0)) return 1.0;
	double sim = dot / den;
	if (sim > 1.0) sim = 1.0;
	if (sim < -1.0) sim = -1.0;
	return 1.0 - sim;
}

/* ------------------------------------------------------------------ bounded max-heap on (dist, label) */

typedef struct {
	double d;
	int64_t id;
} ent_t;

static inline int ent_less(ent_t a, ent_t b) { /* a ranks before b */
	if (a.d < b.d) return 1;
	if (a.d > b.d) return 0;
	/* NaN sorts last */
	if (a.d != a.d) return 0;
	if (b.d != b.d) return 1;
	return a.id < b.id;
}

typedef struct {
	ent_t *h;
	int n, cap;
} heap_t;

static void heap_init(heap_t *hp, int cap) {
	hp->h = (ent_t *)malloc(sizeof(ent_t) * (size_t)(cap > 0 ? cap : 1));
	hp->n = 0;
	hp->cap = cap;
}
static void heap_free(heap_t *hp) { free(hp->h); }

static void heap_sift_down(ent_t *h, int n, int i) {
	for (;;) {
		int l = 2 * i + 1, r = l + 1, m = i;
		if (l < n && ent_less(h[m], h[l])) m = l;
		if (r < n && ent_less(h[m], h[r])) m = r;
		if (m == i) return;
		ent_t t = h[i];
		h[i] = h[m];
		h[m] = t;
		i = m;
	}
}
static void heap_push(heap_t *hp, ent_t e) {
	if (hp->cap <= 0) return;
	if (hp->n < hp->cap) {
		int i = hp->n++;
		hp->h[i] = e;
		while (i > 0) {
			int p = (i - 1) / 2;
			if (!ent_less(hp->h[p], hp->h[i])) break;
			ent_t t = hp->h[p];
			hp->h[p] = hp->h[i];
			hp->h[i] = t;
			i = p;
		}
	} else if (ent_less(e, hp->h[0])) {
		hp->h[0] = e;
		heap_sift_down(hp->h, hp->n, 0);
	}
}
static int ent_cmp(const void *a, const void *b) {
	ent_t x = *(const ent_t *)a, y = *(const ent_t *)b;
	if (ent_less(x, y)) return -1;
	if (ent_less(y, x)) return 1;
	return 0;
}
static void heap_sort_asc(heap_t *hp) { qsort(hp->h, (size_t)hp->n, sizeof(ent_t), ent_cmp); }

static inline int bit_test(const uint32_t *bm, int64_t i) { return (bm[i >> 5] >> (i & 31)) & 1u; }
static inline int row_live(const uint32_t *allow, const uint32_t *deleted, int64_t i) {
	if (deleted && bit_test(deleted, i)) return 0;
	if (allow && !bit_test(allow, i)) return 0;
	return 1;
}

/* ------------------------------------------------------------------ flat search */

int oracle_flat_search(const float *X, int64_t n, int d, int metric, const float *q, int k,
                       const uint32_t *allow, const uint32_t *deleted, int64_t label_base,
                       int64_t *out_labels, float *out_dist32, double *out_dist64) {
	if (k <= 0 || n <= 0) return 0;
	heap_t hp;
	heap_init(&hp, k);
	for (int64_t i = 0; i < n; i++) {
		if (!row_live(allow, deleted, i)) continue;
		ent_t e = {oracle_distance_f64(metric, q, X + i * (int64_t)d, d), label_base + i};
		heap_push(&hp, e);
	}
	heap_sort_asc(&hp);
	int cnt = hp.n;
	for (int i = 0; i < cnt; i++) {
		if (out_labels) out_labels[i] = hp.h[i].id;
		if (out_dist32) out_dist32[i] = (float)hp.h[i].d;
		if (out_dist64) out_dist64[i] = hp.h[i].d;
	}
	heap_free(&hp);
	return cnt;
}

/* fp32, 4 queries per row pass; the compiler vectorises the reductions. */
static inline void dist4_f32(int metric, const float *x, const float *q0, const float *q1,
                             const float *q2, const float *q3, int d, const float *qn,
                             float out[4]) {
	if (metric == ORACLE_L2) {
		float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma omp simd reduction(+ : a0, a1, a2, a3)
		for (int i = 0; i < d; i++) {
			float xi = x[i];
			float e0 = q0[i] - xi, e1 = q1[i] - xi, e2 = q2[i] - xi, e3 = q3[i] - xi;
			a0 += e0 * e0;
			a1 += e1 * e1;
			a2 += e2 * e2;
			a3 += e3 * e3;
		}
		out[0] = a0, out[1] = a1, out[2] = a2, out[3] = a3;
		return;
	}
	float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, nx = 0.f;
#pragma omp simd reduction(+ : a0, a1, a2, a3, nx)
	for (int i = 0; i < d; i++) {
		float xi = x[i];
		a0 += q0[i] * xi;
		a1 += q1[i] * xi;
		a2 += q2[i] * xi;
		a3 += q3[i] * xi;
		nx += xi * xi;
	}
	float a[4] = {a0, a1, a2, a3};
	for (int j = 0; j < 4; j++) {
		if (metric == ORACLE_DOT) {
			out[j] = 1.0f - a[j];
		} else {
			float den = sqrtf(qn[j]) * sqrtf(nx);
			float sim = den > 0.f ? a[j] / den : 0.f;
			if (sim > 1.f) sim = 1.f;
			if (sim < -1.f) sim = -1.f;
			out[j] = 1.0f - sim;
		}
	}
}

int oracle_flat_search_batch_f32(const float *X, int64_t n, int d, int metric, const float *Q,
                                 int nq, int k, const uint32_t *allow, const uint32_t *deleted,
                                 int64_t *out_labels, float *out_dist, int32_t *out_counts,
                                 int nthreads) {
	if (nq <= 0) return 0;
	if (k <= 0 || n <= 0) {
		for (int j = 0; j < nq; j++)
			if (out_counts) out_counts[j] = 0;
		return 0;
	}
#ifdef _OPENMP
	if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
	nthreads = 1;
#endif
	if ((int64_t)nthreads > n) nthreads = (int)n;
	float *qn = (float *)calloc((size_t)nq + 4, sizeof(float));
	for (int j = 0; j < nq; j++) {
		float s = 0.f;
		for (int i = 0; i < d; i++) s += Q[(int64_t)j * d + i] * Q[(int64_t)j * d + i];
		qn[j] = s;
	}
	heap_t *heaps = (heap_t *)malloc(sizeof(heap_t) * (size_t)nthreads * (size_t)nq);
	for (int64_t i = 0; i < (int64_t)nthreads * nq; i++) heap_init(&heaps[i], k);

#pragma omp parallel num_threads(nthreads)
	{
#ifdef _OPENMP
		int t = omp_get_thread_num();
		int T = omp_get_num_threads();
#else
		int t = 0, T = 1;
#endif
		int64_t lo = n * t / T, hi = n * (t + 1) / T;
		heap_t *mine = heaps + (int64_t)t * nq;
		for (int64_t i = lo; i < hi; i++) {
			if (!row_live(allow, deleted, i)) continue;
			const float *x = X + i * (int64_t)d;
			int j = 0;
			for (; j + 4 <= nq; j += 4) {
				float o[4];
				dist4_f32(metric, x, Q + (int64_t)j * d, Q + (int64_t)(j + 1) * d,
				          Q + (int64_t)(j + 2) * d, Q + (int64_t)(j + 3) * d, d, qn + j, o);
				for (int u = 0; u < 4; u++) {
					ent_t e = {(double)o[u], i};
					heap_push(&mine[j + u], e);
				}
			}
			for (; j < nq; j++) {
				const float *q = Q + (int64_t)j * d;
				float o[4];
				dist4_f32(metric, x, q, q, q, q, d, (float[4]){qn[j], qn[j], qn[j], qn[j]}, o);
				ent_t e = {(double)o[0], i};
				heap_push(&mine[j], e);
			}
		}
	}
	/* merge per-thread heaps */
	for (int j = 0; j < nq; j++) {
		heap_t fin;
		heap_init(&fin, k);
		for (int t = 0; t < nthreads; t++) {
			heap_t *h = &heaps[(int64_t)t * nq + j];
			for (int i = 0; i < h->n; i++) heap_push(&fin, h->h[i]);
		}
		heap_sort_asc(&fin);
		for (int i = 0; i < fin.n; i++) {
			out_labels[(int64_t)j * k + i] = fin.h[i].id;
			out_dist[(int64_t)j * k + i] = (float)fin.h[i].d;
		}
		if (out_counts) out_counts[j] = fin.n;
		heap_free(&fin);
	}
	for (int64_t i = 0; i < (int64_t)nthreads * nq; i++) heap_free(&heaps[i]);
	free(heaps);
	free(qn);
	return nq;
}

int oracle_merge_topk(const int64_t *labels, const float *dists, const int32_t *counts, int G,
                      int k, int64_t *out_labels, float *out_dists) {
	heap_t hp;
	heap_init(&hp, k);
	for (int g = 0; g < G; g++)
		for (int i = 0; i < counts[g]; i++) {
			ent_t e = {(double)dists[(int64_t)g * k + i], labels[(int64_t)g * k + i]};
			heap_push(&hp, e);
		}
	heap_sort_asc(&hp);
	int n = hp.n;
	for (int i = 0; i < n; i++) {
		out_labels[i] = hp.h[i].id;
		out_dists[i] = (float)hp.h[i].d;
	}
	heap_free(&hp);
	return n;
}

/* ------------------------------------------------------------------ IVF_PQ */

static void normalise_to(const float *v, int d, float *out) {
	double s = 0.0;
	for (int i = 0; i < d; i++) s += (double)v[i] * (double)v[i];
	double r = s > 0.0 ? 1.0 / sqrt(s) : 1.0;
	for (int i = 0; i < d; i++) out[i] = (float)((double)v[i] * r);
}

/* distance used inside the index: cosine rows are normalised and compared with L2 */
static double index_space_dist(int metric, const float *q, const float *c, int d) {
	if (metric == ORACLE_DOT) return oracle_distance_f64(ORACLE_DOT, q, c, d);
	return oracle_distance_f64(ORACLE_L2, q, c, d);
}

int oracle_ivf_probe(const oracle_ivfpq_t *ix, const float *q, int nprobes, int32_t *out_parts,
                     float *out_dists) {
	int d = ix->dim;
	float *qq = (float *)malloc(sizeof(float) * (size_t)d);
	if (ix->metric == ORACLE_COSINE)
		normalise_to(q, d, qq);
	else
		memcpy(qq, q, sizeof(float) * (size_t)d);
	if (nprobes > ix->P) nprobes = ix->P;
	heap_t hp;
	heap_init(&hp, nprobes);
	for (int p = 0; p < ix->P; p++) {
		ent_t e = {index_space_dist(ix->metric, qq, ix->centroids + (int64_t)p * d, d), p};
		heap_push(&hp, e);
	}
	heap_sort_asc(&hp);
	int n = hp.n;
	for (int i = 0; i < n; i++) {
		out_parts[i] = (int32_t)hp.h[i].id;
		if (out_dists) out_dists[i] = (float)hp.h[i].d;
	}
	heap_free(&hp);
	free(qq);
	return n;
}

void oracle_pq_lut(const oracle_ivfpq_t *ix, const float *q, int part, float *lut) {
	int d = ix->dim, M = ix->M, dsub = d / M;
	float *r = (float *)malloc(sizeof(float) * (size_t)d);
	if (ix->metric == ORACLE_COSINE)
		normalise_to(q, d, r);
	else
		memcpy(r, q, sizeof(float) * (size_t)d);
	if (ix->metric != ORACLE_DOT) {
		const float *c = ix->centroids + (int64_t)part * d;
		for (int i = 0; i < d; i++) r[i] = r[i] - c[i];
	}
	for (int m = 0; m < M; m++) {
		const float *rm = r + m * dsub;
		for (int c = 0; c < 256; c++) {
			const float *cb = ix->codebook + ((int64_t)m * 256 + c) * dsub;
			double acc = 0.0;
			if (ix->metric == ORACLE_DOT) {
				for (int i = 0; i < dsub; i++) acc -= (double)rm[i] * (double)cb[i];
			} else {
				for (int i = 0; i < dsub; i++) {
					double e = (double)rm[i] - (double)cb[i];
					acc += e * e;
				}
			}
			lut[m * 256 + c] = (float)acc;
		}
	}
	free(r);
}

static double adc_sum(const oracle_ivfpq_t *ix, const float *lut, const uint8_t *code) {
	double s = 0.0;
	for (int m = 0; m < ix->M; m++) s += (double)lut[m * 256 + code[m]];
	if (ix->metric == ORACLE_COSINE) return 0.5 * s;
	if (ix->metric == ORACLE_DOT) return 1.0 + s;
	return s;
}

float oracle_adc_distance(const oracle_ivfpq_t *ix, const float *lut, const uint8_t *code) {
	return (float)adc_sum(ix, lut, code);
}

static int ivfpq_search_impl(const oracle_ivfpq_t *ix, const float *X, int64_t n, const float *q, int k,
                             int nprobes, int refine_factor, const uint32_t *allow,
                             const uint32_t *deleted, const int32_t *forced_parts, int32_t *out_parts,
                             int64_t *out_cand_labels, float *out_cand_adc, int32_t *out_ncand,
                             int64_t *out_labels, float *out_dist) {
	int d = ix->dim;
	if (k <= 0) return 0;
	if (nprobes < 1) nprobes = 1;
	if (nprobes > ix->P) nprobes = ix->P;
	int rf = refine_factor < 1 ? 1 : refine_factor;
	int64_t ncand_cap = (int64_t)k * rf;
	int32_t *parts = (int32_t *)malloc(sizeof(int32_t) * (size_t)nprobes);
	int np = nprobes;
	if (forced_parts)
		memcpy(parts, forced_parts, sizeof(int32_t) * (size_t)nprobes);
	else
		np = oracle_ivf_probe(ix, q, nprobes, parts, NULL);
	if (out_parts) memcpy(out_parts, parts, sizeof(int32_t) * (size_t)np);

	float *lut = (float *)malloc(sizeof(float) * (size_t)ix->M * 256);
	heap_t cand;
	heap_init(&cand, (int)ncand_cap);
	for (int pi = 0; pi < np; pi++) {
		int p = parts[pi];
		oracle_pq_lut(ix, q, p, lut);
		for (int64_t r = ix->part_offsets[p]; r < ix->part_offsets[p + 1]; r++) {
			int64_t id = ix->row_ids[r];
			if (!row_live(allow, deleted, id)) continue;
			ent_t e = {(double)(float)adc_sum(ix, lut, ix->codes + r * ix->M), id};
			heap_push(&cand, e);
		}
	}
	heap_sort_asc(&cand);
	if (out_ncand) *out_ncand = cand.n;
	for (int i = 0; i < cand.n; i++) {
		if (out_cand_labels) out_cand_labels[i] = cand.h[i].id;
		if (out_cand_adc) out_cand_adc[i] = (float)cand.h[i].d;
	}

	heap_t fin;
	heap_init(&fin, k);
	for (int i = 0; i < cand.n; i++) {
		ent_t e = cand.h[i];
		if (refine_factor >= 1) e.d = oracle_distance_f64(ix->metric, q, X + e.id * (int64_t)d, d);
		heap_push(&fin, e);
	}
	/* rows appended after the index was built: exact scan, merged (Lance's unindexed fragments) */
	for (int64_t i = ix->n_indexed; i < n; i++) {
		if (!row_live(allow, deleted, i)) continue;
		ent_t e = {oracle_distance_f64(ix->metric, q, X + i * (int64_t)d, d), i};
		heap_push(&fin, e);
	}
	heap_sort_asc(&fin);
	int cnt = fin.n;
	for (int i = 0; i < cnt; i++) {
		out_labels[i] = fin.h[i].id;
		out_dist[i] = (float)fin.h[i].d;
	}
	heap_free(&fin);
	heap_free(&cand);
	free(lut);
	free(parts);
	return cnt;
}

int oracle_ivfpq_search(const oracle_ivfpq_t *ix, const float *X, int64_t n, const float *q, int k,
                        int nprobes, int refine_factor, const uint32_t *allow,
                        const uint32_t *deleted, int32_t *out_parts, int64_t *out_cand_labels,
                        float *out_cand_adc, int32_t *out_ncand, int64_t *out_labels,
                        float *out_dist) {
	return ivfpq_search_impl(ix, X, n, q, k, nprobes, refine_factor, allow, deleted, NULL, out_parts,
	                         out_cand_labels, out_cand_adc, out_ncand, out_labels, out_dist);
}

int oracle_ivfpq_search_parts(const oracle_ivfpq_t *ix, const float *X, int64_t n, const float *q, int k,
                              const int32_t *parts, int nparts, int refine_factor,
                              const uint32_t *allow, const uint32_t *deleted, int64_t *out_cand_labels,
                              float *out_cand_adc, int32_t *out_ncand, int64_t *out_labels,
                              float *out_dist) {
	if (nparts < 1 || nparts > ix->P) return -1;
	return ivfpq_search_impl(ix, X, n, q, k, nparts, refine_factor, allow, deleted, parts, NULL,
	                         out_cand_labels, out_cand_adc, out_ncand, out_labels, out_dist);
}

/* ------------------------------------------------------------------ training */

void oracle_kmeans(const float *X, int64_t n, int d, int k, int iters, int metric, float *cen,
                   int32_t *assign_out) {
	int32_t *assign = assign_out ? assign_out : (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
	for (int c = 0; c < k; c++) {
		int64_t src = n > 0 ? (int64_t)c * n / k : 0;
		if (n > 0)
			memcpy(cen + (int64_t)c * d, X + src * d, sizeof(float) * (size_t)d);
		else
			memset(cen + (int64_t)c * d, 0, sizeof(float) * (size_t)d);
	}
	double *sum = (double *)malloc(sizeof(double) * (size_t)k * (size_t)d);
	int64_t *cnt = (int64_t *)malloc(sizeof(int64_t) * (size_t)k);
	for (int it = 0; it <= iters; it++) {
		/* assignment (the last pass only assigns, so `assign` matches the final centres) */
#pragma omp parallel for schedule(static)
		for (int64_t i = 0; i < n; i++) {
			const float *x = X + i * d;
			double best = DBL_MAX;
			int bi = 0;
			for (int c = 0; c < k; c++) {
				double dist = index_space_dist(metric, x, cen + (int64_t)c * d, d);
				if (dist < best) {
					best = dist;
					bi = c;
				}
			}
			assign[i] = bi;
		}
		if (it == iters) break;
		memset(sum, 0, sizeof(double) * (size_t)k * (size_t)d);
		memset(cnt, 0, sizeof(int64_t) * (size_t)k);
		for (int64_t i = 0; i < n; i++) {
			double *s = sum + (int64_t)assign[i] * d;
			const float *x = X + i * d;
			for (int j = 0; j < d; j++) s[j] += (double)x[j];
			cnt[assign[i]]++;
		}
		for (int c = 0; c < k; c++) {
			if (cnt[c] == 0) continue;
			for (int j = 0; j < d; j++) cen[(int64_t)c * d + j] = (float)(sum[(int64_t)c * d + j] / (double)cnt[c]);
		}
	}
	free(sum);
	free(cnt);
	if (!assign_out) free(assign);
}

void oracle_ivfpq_train(const float *X, int64_t n, int d, int metric, int P, int M, int iters,
                        float *centroids, float *codebook, int64_t *part_offsets,
                        int64_t *row_ids, uint8_t *codes) {
	int dsub = d / M;
	/* cosine: the index lives on normalised rows */
	float *Xn = (float *)malloc(sizeof(float) * (size_t)n * (size_t)d);
	for (int64_t i = 0; i < n; i++) {
		if (metric == ORACLE_COSINE)
			normalise_to(X + i * d, d, Xn + i * d);
		else
			memcpy(Xn + i * d, X + i * d, sizeof(float) * (size_t)d);
	}
	int32_t *assign = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
	oracle_kmeans(Xn, n, d, P, iters, metric, centroids, assign);

	/* residuals (L2 / cosine); dot keeps the raw vectors */
	float *R = (float *)malloc(sizeof(float) * (size_t)n * (size_t)d);
	for (int64_t i = 0; i < n; i++) {
		const float *c = centroids + (int64_t)assign[i] * d;
		for (int j = 0; j < d; j++) R[i * d + j] = metric == ORACLE_DOT ? Xn[i * d + j] : Xn[i * d + j] - c[j];
	}
	/* per-subspace codebooks */
	float *sub = (float *)malloc(sizeof(float) * (size_t)n * (size_t)dsub);
	int32_t *sassign = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
	uint8_t *codes_rowmajor = (uint8_t *)malloc((size_t)n * (size_t)M);
	for (int m = 0; m < M; m++) {
		for (int64_t i = 0; i < n; i++) memcpy(sub + i * dsub, R + i * d + m * dsub, sizeof(float) * (size_t)dsub);
		oracle_kmeans(sub, n, dsub, 256, iters, ORACLE_L2, codebook + (int64_t)m * 256 * dsub, sassign);
		for (int64_t i = 0; i < n; i++) codes_rowmajor[i * M + m] = (uint8_t)sassign[i];
	}
	/* partition-major order, labels ascending inside a partition */
	memset(part_offsets, 0, sizeof(int64_t) * (size_t)(P + 1));
	for (int64_t i = 0; i < n; i++) part_offsets[assign[i] + 1]++;
	for (int p = 0; p < P; p++) part_offsets[p + 1] += part_offsets[p];
	int64_t *cursor = (int64_t *)malloc(sizeof(int64_t) * (size_t)P);
	memcpy(cursor, part_offsets, sizeof(int64_t) * (size_t)P);
	for (int64_t i = 0; i < n; i++) {
		int64_t pos = cursor[assign[i]]++;
		row_ids[pos] = i;
		memcpy(codes + pos * M, codes_rowmajor + i * M, (size_t)M);
	}
	free(cursor);
	free(codes_rowmajor);
	free(sassign);
	free(sub);
	free(R);
	free(assign);
	free(Xn);
}

/* ------------------------------------------------------------------ chunked fp64 accumulator
 * Same ranking truth as oracle_flat_search (fp64 distance, ties by label), but the table arrives in
 * row chunks and several queries share one pass over each chunk, OpenMP over the rows of a chunk.
 * This is what lets the full-size configurations of BASELINE.json (100 M x 128, 10 M x 384 ...) be
 * checked against the oracle without ever holding the table on the host. */
struct oracle_acc {
	int metric, nq, d, k;
	double *Qd;  /* [nq][d] */
	double *qn;  /* [nq] |q|^2 */
	heap_t *heaps; /* [nq] */
};

oracle_acc_t *oracle_acc_new(int metric, const float *Q, int nq, int d, int k) {
	oracle_acc_t *a = (oracle_acc_t *)calloc(1, sizeof(oracle_acc_t));
	a->metric = metric, a->nq = nq, a->d = d, a->k = k;
	a->Qd = (double *)malloc(sizeof(double) * (size_t)nq * (size_t)d);
	a->qn = (double *)calloc((size_t)nq, sizeof(double));
	a->heaps = (heap_t *)malloc(sizeof(heap_t) * (size_t)nq);
	for (int j = 0; j < nq; j++) {
		heap_init(&a->heaps[j], k);
		for (int i = 0; i < d; i++) {
			double v = (double)Q[(int64_t)j * d + i];
			a->Qd[(int64_t)j * d + i] = v;
			a->qn[j] += v * v;
		}
	}
	return a;
}

void oracle_acc_add(oracle_acc_t *a, const float *X, int64_t n, int64_t row0, const uint32_t *allow,
                    int64_t allow_bits, const uint32_t *deleted, int64_t label_base, int nthreads) {
	if (n <= 0 || a->k <= 0) return;
#ifdef _OPENMP
	if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
	nthreads = 1;
#endif
	if ((int64_t)nthreads > n) nthreads = (int)n;
	const int nq = a->nq, d = a->d, metric = a->metric;
	heap_t *heaps = (heap_t *)malloc(sizeof(heap_t) * (size_t)nthreads * (size_t)nq);
	for (int64_t i = 0; i < (int64_t)nthreads * nq; i++) heap_init(&heaps[i], a->k);
#pragma omp parallel num_threads(nthreads)
	{
#ifdef _OPENMP
		int t = omp_get_thread_num(), T = omp_get_num_threads();
#else
		int t = 0, T = 1;
#endif
		int64_t lo = n * t / T, hi = n * (t + 1) / T;
		heap_t *mine = heaps + (int64_t)t * nq;
		double *xd = (double *)malloc(sizeof(double) * (size_t)d);
		for (int64_t i = lo; i < hi; i++) {
			const int64_t row = row0 + i;
			if (deleted && bit_test(deleted, row)) continue;
			if (allow && !(row < allow_bits && bit_test(allow, row))) continue;
			const float *x = X + i * (int64_t)d;
			double nx = 0.0;
			for (int c = 0; c < d; c++) {
				xd[c] = (double)x[c];
				nx += xd[c] * xd[c];
			}
			for (int j = 0; j < nq; j++) {
				const double *q = a->Qd + (int64_t)j * d;
				double acc = 0.0, dist;
				if (metric == ORACLE_L2) {
#pragma omp simd reduction(+ : acc)
					for (int c = 0; c < d; c++) {
						double e = q[c] - xd[c];
						acc += e * e;
					}
					dist = acc;
				} else {
#pragma omp simd reduction(+ : acc)
					for (int c = 0; c < d; c++) acc += q[c] * xd[c];
					if (metric == ORACLE_DOT)
						dist = 1.0 - acc;
					else {
						double den = sqrt(a->qn[j]) * sqrt(nx);
						if (!(den > 0.0))
							dist = 1.0;
						else {
							double sim = acc / den;
							if (sim > 1.0) sim = 1.0;
							if (sim < -1.0) sim = -1.0;
							dist = 1.0 - sim;
						}
					}
				}
				ent_t e = {dist, label_base + row};
				heap_push(&mine[j], e);
			}
		}
		free(xd);
	}
	for (int j = 0; j < nq; j++)
		for (int t = 0; t < nthreads; t++) {
			heap_t *h = &heaps[(int64_t)t * nq + j];
			for (int i = 0; i < h->n; i++) heap_push(&a->heaps[j], h->h[i]);
		}
	for (int64_t i = 0; i < (int64_t)nthreads * nq; i++) heap_free(&heaps[i]);
	free(heaps);
}

void oracle_acc_finish(oracle_acc_t *a, int64_t *out_labels, double *out_dist64, int32_t *out_counts) {
	for (int j = 0; j < a->nq; j++) {
		heap_t *h = &a->heaps[j];
		heap_sort_asc(h);
		for (int i = 0; i < h->n; i++) {
			out_labels[(int64_t)j * a->k + i] = h->h[i].id;
			out_dist64[(int64_t)j * a->k + i] = h->h[i].d;
		}
		out_counts[j] = h->n;
		heap_free(h);
	}
	free(a->heaps);
	free(a->Qd);
	free(a->qn);
	free(a);
}

/* ------------------------------------------------------------------ blocked fp32 batch search
 * The timed CPU baseline SURVEY.md 8(d) asks for: "blocked SGEMM-style flat search + heap top-k, all
 * host cores".  Queries are packed into panels of 16 (k-major, so one AVX2 register pair holds the 16
 * queries' element i), rows are taken 6 at a time: a 6 x 16 register tile (12 accumulators), every
 * row element broadcast once per panel.  L2 uses |q|^2 + |x|^2 - 2 q.x, cosine the same dot product.
 * Row blocks are sized to stay cache resident while all the panels pass over them.  Ranking rule as
 * everywhere: ascending (distance, label). */
typedef float v8f __attribute__((vector_size(32), aligned(4)));
#define OB_QP 16
#define OB_RT 6

static inline void micro_6x16(const float *const x[OB_RT], const float *qt, int d, float out[OB_RT][OB_QP]) {
	v8f a00 = {0}, a01 = {0}, a10 = {0}, a11 = {0}, a20 = {0}, a21 = {0};
	v8f a30 = {0}, a31 = {0}, a40 = {0}, a41 = {0}, a50 = {0}, a51 = {0};
	const float *x0 = x[0], *x1 = x[1], *x2 = x[2], *x3 = x[3], *x4 = x[4], *x5 = x[5];
	for (int i = 0; i < d; i++) {
		const v8f q0 = *(const v8f *)(qt + (int64_t)i * OB_QP), q1 = *(const v8f *)(qt + (int64_t)i * OB_QP + 8);
		a00 += x0[i] * q0, a01 += x0[i] * q1;
		a10 += x1[i] * q0, a11 += x1[i] * q1;
		a20 += x2[i] * q0, a21 += x2[i] * q1;
		a30 += x3[i] * q0, a31 += x3[i] * q1;
		a40 += x4[i] * q0, a41 += x4[i] * q1;
		a50 += x5[i] * q0, a51 += x5[i] * q1;
	}
	*(v8f *)&out[0][0] = a00, *(v8f *)&out[0][8] = a01;
	*(v8f *)&out[1][0] = a10, *(v8f *)&out[1][8] = a11;
	*(v8f *)&out[2][0] = a20, *(v8f *)&out[2][8] = a21;
	*(v8f *)&out[3][0] = a30, *(v8f *)&out[3][8] = a31;
	*(v8f *)&out[4][0] = a40, *(v8f *)&out[4][8] = a41;
	*(v8f *)&out[5][0] = a50, *(v8f *)&out[5][8] = a51;
}

int oracle_flat_search_batch_blocked(const float *X, int64_t n, int d, int metric, const float *Q, int nq,
                                     int k, const uint32_t *allow, const uint32_t *deleted,
                                     int64_t *out_labels, float *out_dist, int32_t *out_counts,
                                     int nthreads) {
	if (nq <= 0) return 0;
	if (k <= 0 || n <= 0) {
		for (int j = 0; j < nq; j++)
			if (out_counts) out_counts[j] = 0;
		return 0;
	}
#ifdef _OPENMP
	if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
	nthreads = 1;
#endif
	if ((int64_t)nthreads > n) nthreads = (int)n;
	const int npan = (nq + OB_QP - 1) / OB_QP;
	const int nqp = npan * OB_QP;
	/* panels: Qt[p][i][16] */
	float *Qt = (float *)aligned_alloc(64, sizeof(float) * (size_t)npan * (size_t)d * OB_QP);
	float *qn = (float *)calloc((size_t)nqp, sizeof(float));
	for (int p = 0; p < npan; p++)
		for (int i = 0; i < d; i++)
			for (int u = 0; u < OB_QP; u++) {
				int j = p * OB_QP + u;
				Qt[((int64_t)p * d + i) * OB_QP + u] = j < nq ? Q[(int64_t)j * d + i] : 0.f;
			}
	for (int j = 0; j < nq; j++) {
		float s = 0.f;
		for (int i = 0; i < d; i++) s += Q[(int64_t)j * d + i] * Q[(int64_t)j * d + i];
		qn[j] = s;
	}
	int RB = (int)(262144 / ((int64_t)d * 4)); /* ~256 KB of rows per block */
	if (RB < 2 * OB_RT) RB = 2 * OB_RT;
	if (RB > 1024) RB = 1024;
	heap_t *heaps = (heap_t *)malloc(sizeof(heap_t) * (size_t)nthreads * (size_t)nq);
	for (int64_t i = 0; i < (int64_t)nthreads * nq; i++) heap_init(&heaps[i], k);

#pragma omp parallel num_threads(nthreads)
	{
#ifdef _OPENMP
		int t = omp_get_thread_num(), T = omp_get_num_threads();
#else
		int t = 0, T = 1;
#endif
		int64_t lo = n * t / T, hi = n * (t + 1) / T;
		heap_t *mine = heaps + (int64_t)t * nq;
		float *thr = (float *)malloc(sizeof(float) * (size_t)nqp);
		for (int j = 0; j < nqp; j++) thr[j] = j < nq ? INFINITY : -INFINITY;
		const float **rows = (const float **)malloc(sizeof(float *) * (size_t)(RB + OB_RT));
		int64_t *ids = (int64_t *)malloc(sizeof(int64_t) * (size_t)(RB + OB_RT));
		float *xn = (float *)malloc(sizeof(float) * (size_t)(RB + OB_RT));
		float out[OB_RT][OB_QP] __attribute__((aligned(32)));
		for (int64_t b0 = lo; b0 < hi; b0 += RB) {
			int64_t b1 = b0 + RB < hi ? b0 + RB : hi;
			int m = 0;
			for (int64_t i = b0; i < b1; i++) {
				if (!row_live(allow, deleted, i)) continue;
				rows[m] = X + i * (int64_t)d;
				ids[m] = i;
				m++;
			}
			if (m == 0) continue;
			if (metric != ORACLE_DOT)
				for (int r = 0; r < m; r++) {
					float s = 0.f;
					const float *x = rows[r];
#pragma omp simd reduction(+ : s)
					for (int i = 0; i < d; i++) s += x[i] * x[i];
					xn[r] = s;
				}
			for (int r = m; r < m + OB_RT; r++) rows[r] = rows[m - 1]; /* padding of the last tile */
			for (int p = 0; p < npan; p++) {
				const float *qt = Qt + (int64_t)p * d * OB_QP;
				const float *tp = thr + p * OB_QP;
				for (int r0 = 0; r0 < m; r0 += OB_RT) {
					micro_6x16(rows + r0, qt, d, out);
					const int rr = m - r0 < OB_RT ? m - r0 : OB_RT;
					for (int r = 0; r < rr; r++) {
						for (int u = 0; u < OB_QP; u++) {
							float dot = out[r][u], dist;
							const int j = p * OB_QP + u;
							if (metric == ORACLE_DOT)
								dist = 1.0f - dot;
							else if (metric == ORACLE_L2) {
								dist = qn[j] + xn[r0 + r] - 2.0f * dot;
								if (dist < 0.f) dist = 0.f;
							} else {
								float den = sqrtf(qn[j]) * sqrtf(xn[r0 + r]);
								float sim = den > 0.f ? dot / den : 0.f;
								if (sim > 1.f) sim = 1.f;
								if (sim < -1.f) sim = -1.f;
								dist = 1.0f - sim;
							}
							if (dist <= tp[u]) {
								ent_t e = {(double)dist, ids[r0 + r]};
								heap_push(&mine[j], e);
								if (mine[j].n == mine[j].cap) thr[j] = (float)mine[j].h[0].d;
							}
						}
					}
				}
			}
		}
		free(xn);
		free(ids);
		free(rows);
		free(thr);
	}
	for (int j = 0; j < nq; j++) {
		heap_t fin;
		heap_init(&fin, k);
		for (int t = 0; t < nthreads; t++) {
			heap_t *h = &heaps[(int64_t)t * nq + j];
			for (int i = 0; i < h->n; i++) heap_push(&fin, h->h[i]);
		}
		heap_sort_asc(&fin);
		for (int i = 0; i < fin.n; i++) {
			out_labels[(int64_t)j * k + i] = fin.h[i].id;
			out_dist[(int64_t)j * k + i] = (float)fin.h[i].d;
		}
		if (out_counts) out_counts[j] = fin.n;
		heap_free(&fin);
	}
	for (int64_t i = 0; i < (int64_t)nthreads * nq; i++) heap_free(&heaps[i]);
	free(heaps);
	free(qn);
	free(Qt);
	return nq;
}
