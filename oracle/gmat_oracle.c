/*
 * TEST INFRASTRUCTURE ONLY -- plain C restatement of the reference's native hot path.
 * Nothing under gmat_b200/ may link or call this file; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg use it (as the checker / timed CPU baseline, never as the product).
 *
 * Written from the behaviour of (never copied from):
 *   gmat/process_plink/_read_plink_bed.c:19-46    byte walk, 4 samples per byte low bits first,
 *                                                 value (c*c+c)/6.0, short last byte per SNP
 *   gmat/remma/_remma_epi_eff_cpu.c:96-112        additive centring, p accumulated as sum g/(2n)
 *   gmat/remma/_remma_epi_eff_cpu.c:272-287       A and D coding for the AD screen (|g-2|<1e-4 -> 0)
 *   gmat/remma/_remma_epi_eff_cpu.c:454-466       D coding for the DD screen
 *   gmat/remma/_remma_epi_eff_cpu.c:69-77         eff = sum_k x_i x_j py, strict '>' threshold
 *   gmat/remma/_remma_epi_eff_cpu.c:233-247       AD: (i,j) with '>=' then (j,i) with '>'
 *   gmat/remma/_remma_epi_eff_cpu.c:154,332,339   MAF-binned threshold table index
 *   gmat/remma/remma_epiAA/remma_epiAA.py:72-77   exact statistics eff, var = w'Aw, chi
 * Parity of this port is pinned by tests/test_oracle.py against oracle/_ref (the reference C itself,
 * compiled by oracle/Makefile) and against tests/golden (outputs of the reference run here).
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* out[snp*n + k] = (c*c+c)/6.0; returns 1, or -2 when the file cannot be read */
int oracle_read_bed(const char *prefix, long long n, long long m, double *out) {
    char path[4096];
    snprintf(path, sizeof(path), "%s.bed", prefix);
    FILE *f = fopen(path, "rb");
    if (!f) return -2;
    long long bytes = (n + 3) / 4;
    unsigned char *row = (unsigned char *)malloc((size_t)bytes);
    if (fseek(f, 3, SEEK_SET) != 0) { fclose(f); free(row); return -2; }
    for (long long s = 0; s < m; ++s) {
        if ((long long)fread(row, 1, (size_t)bytes, f) != bytes) { fclose(f); free(row); return -2; }
        for (long long k = 0; k < n; ++k) {
            int c = (row[k >> 2] >> (2 * (k & 3))) & 3;
            out[s * n + k] = (c * c + c) / 6.0;
        }
    }
    free(row);
    fclose(f);
    return 1;
}

/* in place: raw -> coded.  coding 1: g - 2p ; 2: h - 2p(1-p).  p summed term by term like the C scan. */
void oracle_code(double *x, long long n, long long m, int coding, double *freq_out) {
    for (long long s = 0; s < m; ++s) {
        double p = 0.0;
        double *r = x + s * n;
        for (long long k = 0; k < n; ++k) p += r[k] / (2 * n);
        if (freq_out) freq_out[s] = p;
        for (long long k = 0; k < n; ++k) {
            if (coding == 1) {
                r[k] -= 2 * p;
            } else {
                double h = (fabs(r[k] - 2.0) < 0.0001) ? 0.0 : r[k];
                r[k] = h - 2 * p * (1 - p);
            }
        }
    }
}

/* generic screen: for i in ilist, j in (i, m): eff = sum_k U[i][k] V[j][k] py[k].
 * cut: scalar when table == NULL, else table[binI[i]*10 + binJ[j]].  ge != 0 -> '>='.
 * Hits are appended (i, j, eff) to the caller's arrays up to cap; returns the total count. */
long long oracle_screen(const double *U, const double *V, long long n, long long m, const long long *ilist,
                        long long n_i, const double *py, double cut, const double *table,
                        const long long *binI, const long long *binJ, int ge, long long *hi, long long *hj,
                        double *heff, long long cap) {
    long long count = 0;
#pragma omp parallel for schedule(guided, 5)
    for (long long q = 0; q < n_i; ++q) {
        const long long i = ilist[q];
        const double *u = U + i * n;
        for (long long j = i + 1; j < m; ++j) {
            const double *v = V + j * n;
            double e = 0.0;
            for (long long k = 0; k < n; ++k) e += u[k] * v[k] * py[k];
            const double c = table ? table[binI[i] * 10 + binJ[j]] : cut;
            const int hit = ge ? (fabs(e) >= c) : (fabs(e) > c);
            if (hit) {
                long long pos;
#pragma omp atomic capture
                pos = count++;
                if (pos < cap) { hi[pos] = i; hj[pos] = j; heff[pos] = e; }
            }
        }
    }
    return count;
}

/* exact statistics for explicit pairs: w = U[i] o V[j]; eff = w'py; var = w'Aw (A row-major n x n). */
void oracle_exact_pairs(const double *U, const double *V, long long n, const long long *pi, const long long *pj,
                        long long n_pairs, const double *py, const double *A, double *eff, double *var) {
#pragma omp parallel
    {
        double *w = (double *)malloc(sizeof(double) * (size_t)n);
#pragma omp for schedule(dynamic, 4)
        for (long long q = 0; q < n_pairs; ++q) {
            const double *u = U + pi[q] * n, *v = V + pj[q] * n;
            double e = 0.0, s = 0.0;
            for (long long k = 0; k < n; ++k) { w[k] = u[k] * v[k]; e += w[k] * py[k]; }
            for (long long k = 0; k < n; ++k) {
                const double *a = A + k * n;
                double y = 0.0;
                for (long long l = 0; l < n; ++l) y += a[l] * w[l];
                s += w[k] * y;
            }
            eff[q] = e;
            var[q] = s;
        }
        free(w);
    }
}
