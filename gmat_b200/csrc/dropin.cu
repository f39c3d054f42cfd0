// The seven DROP-IN symbols: same names / signatures / file side effects as the reference's cffi
// modules (gmat/remma/_build.py:8-31, gmat/process_plink/_build.py:8), running on the GPU.
// Host pointers in, text file out.  These own their device memory for the duration of the call.
//
// Deliberate differences from the reference C (documented in INTEGRATION.md):
//   * failure -> negative return + gmat_b200_last_error(), never exit(1)   (_remma_epi_eff_cpu.c:18-21)
//   * the .bed magic is validated                                          (_read_plink_bed.c:31 skips it)
//   * rows are written sorted by (loop SNP, other SNP, direction) instead of in OpenMP race order
//   * no progress bar on stdout
#include "common.cuh"
#include <algorithm>
#include <cerrno>
#include <cstring>
#include <string>
#include <vector>

using namespace gb;

namespace {

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t bytes) {
        if (bytes == 0) bytes = 16;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            set_error("cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
            p = nullptr;
            return GMAT_B200_ERR_CUDA;
        }
        return GMAT_B200_OK;
    }
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

#define GB_TRY(expr)                           \
    do {                                       \
        int rc_ = (expr);                      \
        if (rc_ != GMAT_B200_OK) return rc_;   \
    } while (0)

struct Store {                     // genotype store on the device
    DevBuf bed, counts;
    long long n = 0, m = 0, row_bytes = 0;
    int load(const char *bed_file, long long n_, long long m_) {
        n = n_; m = m_;
        row_bytes = round_up((n + 3) / 4, 32);
        std::vector<unsigned char> host((size_t)(row_bytes * m));
        GB_TRY(gmat_b200_bed_read_host(bed_file, n, m, row_bytes, host.data()));
        GB_TRY(bed.alloc(host.size()));
        GB_TRY(counts.alloc(sizeof(int) * 4 * (size_t)m));
        GB_CUDA(cudaMemcpy(bed.p, host.data(), host.size(), cudaMemcpyHostToDevice));
        GB_TRY(gmat_b200_bed_stats(bed.as<unsigned char>(), n, m, row_bytes, counts.as<int>(), nullptr));
        return GMAT_B200_OK;
    }
    int lut(int coding, DevBuf &out) {
        GB_TRY(out.alloc(sizeof(double) * 4 * (size_t)m));
        // FREQ_CLOOP: p accumulated exactly like the C scan (_remma_epi_eff_cpu.c:103-107)
        return gmat_b200_build_lut(bed.as<unsigned char>(), n, m, row_bytes, counts.as<int>(), coding,
                                   GMAT_B200_FREQ_CLOOP, out.as<double>(), nullptr, nullptr);
    }
};

struct Pass {                      // one kernel pass of the screen
    int codingU, codingV, flags;
    const long long *binI, *binJ;  // host long long arrays or NULL
};

int run_screen(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
               long long len_snp_lst_0, double *pymat, double eff_cut, const double *cut_table,
               const Pass *passes, int n_pass, char *out_file) {
    GB_CHECK_ARG(bed_file && out_file && pymat, "NULL pointer");
    GB_CHECK_ARG(num_id > 0 && num_snp > 0 && len_snp_lst_0 >= 0, "sizes");
    GB_CHECK_ARG(len_snp_lst_0 == 0 || snp_lst_0, "snp_lst_0 is NULL");
    for (long long q = 0; q < len_snp_lst_0; ++q)
        GB_CHECK_ARG(snp_lst_0[q] >= 0 && snp_lst_0[q] < num_snp, "snp_lst_0 entry out of range");
    Store st;
    GB_TRY(st.load(bed_file, num_id, num_snp));
    DevBuf lutA, lutD, py, ilist, hits, count, table, binI, binJ;
    bool needA = false, needD = false;
    for (int q = 0; q < n_pass; ++q) {
        needA |= passes[q].codingU == GMAT_B200_CODING_ADD || passes[q].codingV == GMAT_B200_CODING_ADD;
        needD |= passes[q].codingU == GMAT_B200_CODING_DOM || passes[q].codingV == GMAT_B200_CODING_DOM;
    }
    if (needA) GB_TRY(st.lut(GMAT_B200_CODING_ADD, lutA));
    if (needD) GB_TRY(st.lut(GMAT_B200_CODING_DOM, lutD));
    GB_TRY(py.alloc(sizeof(double) * (size_t)num_id));
    GB_CUDA(cudaMemcpy(py.p, pymat, sizeof(double) * (size_t)num_id, cudaMemcpyHostToDevice));
    GB_TRY(ilist.alloc(sizeof(long long) * (size_t)len_snp_lst_0));
    if (len_snp_lst_0)
        GB_CUDA(cudaMemcpy(ilist.p, snp_lst_0, sizeof(long long) * (size_t)len_snp_lst_0, cudaMemcpyHostToDevice));
    GB_TRY(count.alloc(sizeof(unsigned long long)));
    if (cut_table) {
        GB_TRY(table.alloc(sizeof(double) * 111));
        GB_CUDA(cudaMemcpy(table.p, cut_table, sizeof(double) * 111, cudaMemcpyHostToDevice));
    }

    // engine of the screen: int8 tensor-core prefilter + FP64 re-test (screen_i8.cu; GMAT_B200_SCREEN_ENGINE=f64 forces
    // the DMMA kernel).  Missing calls (code 01 = 1/3 in the C path) select the tripled integer coding.
    const char *f64_only = getenv("GMAT_B200_SCREEN_ENGINE");
    if (f64_only != nullptr && strcmp(f64_only, "f64") != 0) f64_only = nullptr;
    int tripled = 0;
    {
        std::vector<int> hc((size_t)(4 * num_snp));
        GB_CUDA(cudaMemcpy(hc.data(), st.counts.p, sizeof(int) * hc.size(), cudaMemcpyDeviceToHost));
        long long missing = 0;
        for (long long s = 0; s < num_snp; ++s) missing += hc[(size_t)(4 * s + 1)];
        tripled = missing != 0;
    }
    bool i8 = f64_only == nullptr && num_id < (tripled ? 29000 : 60000) && len_snp_lst_0 > 0;
    // operands of the int8 screen, built ONCE for all passes of the call (the two directions of AD share them)
    DevBuf g8A, g8D, tA, tD, dig, stats, cand, scr;
    const long long cand_cap = 1LL << 22;
    if (i8) {
        const long long ld = gmat_b200_screen_ld(num_id);
        if (needA) {
            GB_TRY(g8A.alloc((size_t)num_snp * ld));
            GB_TRY(tA.alloc(sizeof(double) * (size_t)num_snp));
            GB_TRY(gmat_b200_screen_operand(st.bed.as<unsigned char>(), st.n, st.m, st.row_bytes, GMAT_B200_CODING_ADD,
                                            tripled, py.as<double>(), g8A.as<signed char>(), tA.as<double>(), nullptr));
        }
        if (needD) {
            GB_TRY(g8D.alloc((size_t)num_snp * ld));
            GB_TRY(tD.alloc(sizeof(double) * (size_t)num_snp));
            GB_TRY(gmat_b200_screen_operand(st.bed.as<unsigned char>(), st.n, st.m, st.row_bytes, GMAT_B200_CODING_DOM,
                                            tripled, py.as<double>(), g8D.as<signed char>(), tD.as<double>(), nullptr));
        }
        GB_TRY(dig.alloc((size_t)3 * ld));
        GB_TRY(stats.alloc(sizeof(double) * 4));
        GB_TRY(gmat_b200_screen_py(py.as<double>(), st.n, tripled, dig.as<signed char>(), stats.as<double>(), nullptr));
        GB_TRY(cand.alloc(sizeof(int) * 2 * (size_t)cand_cap));
        GB_TRY(scr.alloc(64));                                    // [0] candidate count (u64), [8] status (int)
    }

    std::vector<gmat_b200_hit> all;
    std::vector<int> dir;
    if (i8) set_engine_note("int8 tensor-core prefilter + FP64 re-test (%s)",
                            tripled ? "3 base-32 planes: the store has missing calls" : "2 base-128 planes");
    else set_engine_note("FP64 DMMA kernel (%s)", f64_only ? "GMAT_B200_SCREEN_ENGINE=f64"
                                                            : (len_snp_lst_0 == 0 ? "empty first-SNP list"
                                                                                  : "n beyond the int8 screen's range"));
    for (int q = 0; q < n_pass; ++q) {
        const Pass &ps = passes[q];
        DevBuf bI, bJ;
        if (cut_table) {
            GB_CHECK_ARG(ps.binI && ps.binJ, "freq arrays are NULL");
            std::vector<int> hI((size_t)num_snp), hJ((size_t)num_snp);
            for (long long s = 0; s < num_snp; ++s) {
                GB_CHECK_ARG(ps.binI[s] >= 0 && ps.binI[s] <= 10 && ps.binJ[s] >= 0 && ps.binJ[s] <= 10,
                             "freq bin outside 0..10");
                hI[(size_t)s] = (int)ps.binI[s];
                hJ[(size_t)s] = (int)ps.binJ[s];
            }
            GB_TRY(bI.alloc(sizeof(int) * (size_t)num_snp));
            GB_TRY(bJ.alloc(sizeof(int) * (size_t)num_snp));
            GB_CUDA(cudaMemcpy(bI.p, hI.data(), sizeof(int) * (size_t)num_snp, cudaMemcpyHostToDevice));
            GB_CUDA(cudaMemcpy(bJ.p, hJ.data(), sizeof(int) * (size_t)num_snp, cudaMemcpyHostToDevice));
        }
        const bool uA = ps.codingU == GMAT_B200_CODING_ADD, vA = ps.codingV == GMAT_B200_CODING_ADD;
        const double *lu = uA ? lutA.as<double>() : lutD.as<double>();
        const double *lv = vA ? lutA.as<double>() : lutD.as<double>();
        long long cap = 1 << 20;
        bool use_i8 = i8;
        for (int attempt = 0; attempt < 4; ++attempt) {
            DevBuf h;
            GB_TRY(h.alloc(sizeof(gmat_b200_hit) * (size_t)cap));
            GB_CUDA(cudaMemset(count.p, 0, sizeof(unsigned long long)));
            if (use_i8)
                GB_TRY(gmat_b200_epi_eff_i8p(
                    st.bed.as<unsigned char>(), st.n, st.m, st.row_bytes, lu, lv,
                    uA ? g8A.as<signed char>() : g8D.as<signed char>(), vA ? g8A.as<signed char>() : g8D.as<signed char>(),
                    uA ? tA.as<double>() : tD.as<double>(), vA ? tA.as<double>() : tD.as<double>(),
                    dig.as<signed char>(), stats.as<double>(), tripled, py.as<double>(), ilist.as<long long>(),
                    len_snp_lst_0, ps.flags, eff_cut, cut_table ? table.as<double>() : nullptr, bI.as<int>(), bJ.as<int>(),
                    h.as<gmat_b200_hit>(), cap, count.as<unsigned long long>(), cand.as<int>(), cand_cap,
                    scr.as<unsigned long long>(), scr.as<int>() + 2, nullptr));
            else
                GB_TRY(gmat_b200_epi_eff(st.bed.as<unsigned char>(), st.n, st.m, st.row_bytes, lu, lv,
                                         py.as<double>(), ilist.as<long long>(), len_snp_lst_0, ps.flags, eff_cut,
                                         cut_table ? table.as<double>() : nullptr, bI.as<int>(), bJ.as<int>(),
                                         h.as<gmat_b200_hit>(), cap, count.as<unsigned long long>(), nullptr));
            unsigned long long cnt = 0;
            GB_CUDA(cudaMemcpy(&cnt, count.p, sizeof(cnt), cudaMemcpyDeviceToHost));   // the one synchronisation
            if (use_i8) {
                int status = 0;
                GB_CUDA(cudaMemcpy(&status, scr.as<int>() + 2, sizeof(int), cudaMemcpyDeviceToHost));
                if (status != 0) {                                // the int8 path declined: FP64 DMMA kernel
                    set_engine_note("FP64 DMMA kernel (%s)", status == 1 ? "margin of the int8 prefilter not small "
                                    "against the cut" : "more stage-1 survivors than the candidate buffer");
                    use_i8 = false;
                    continue;
                }
            }
            if ((long long)cnt > cap) { cap = (long long)cnt; continue; }   // exact count known: retry once
            const size_t old = all.size();
            all.resize(old + (size_t)cnt);
            if (cnt) GB_CUDA(cudaMemcpy(all.data() + old, h.p, sizeof(gmat_b200_hit) * (size_t)cnt,
                                        cudaMemcpyDeviceToHost));
            dir.resize(all.size(), q);
            break;
        }
    }
    // deterministic row order: (loop SNP, other SNP, direction) == the single-threaded C order
    std::vector<size_t> order(all.size());
    for (size_t q = 0; q < order.size(); ++q) order[q] = q;
    auto key = [&](size_t q, long long &a, long long &b, int &d) {
        const bool sw = (passes[dir[q]].flags & GMAT_B200_EFF_SWAP) != 0;
        a = sw ? all[q].snp_1 : all[q].snp_0;
        b = sw ? all[q].snp_0 : all[q].snp_1;
        d = dir[q];
    };
    std::sort(order.begin(), order.end(), [&](size_t x, size_t y) {
        long long ax, bx, ay, by; int dx, dy;
        key(x, ax, bx, dx); key(y, ay, by, dy);
        if (ax != ay) return ax < ay;
        if (bx != by) return bx < by;
        return dx < dy;
    });
    FILE *f = fopen(out_file, "w");
    if (!f) {
        set_error("Fail to build the output file %s (%s)", out_file, strerror(errno));
        return GMAT_B200_ERR_IO;
    }
    fprintf(f, "%s %s %s\n", "snp_0", "snp_1", "eff");
    for (size_t q : order)
        fprintf(f, "%lld %lld %g\n", (long long)all[q].snp_0, (long long)all[q].snp_1, all[q].eff);
    if (fclose(f) != 0) {
        set_error("error closing %s (%s)", out_file, strerror(errno));
        return GMAT_B200_ERR_IO;
    }
    return GMAT_B200_OK;
}

}  // namespace

extern "C" int read_plink_bed(char *bed_file, long long num_id, long long num_snp, double *marker_mat) {
    GB_CHECK_ARG(bed_file && marker_mat, "NULL pointer");
    GB_CHECK_ARG(num_id > 0 && num_snp > 0, "sizes");
    Store st;
    GB_TRY(st.load(bed_file, num_id, num_snp));
    DevBuf out;
    // decode in SNP chunks so the device buffer stays bounded (8 bytes per genotype)
    const long long chunk = std::max<long long>(1, (1LL << 28) / num_id);
    GB_TRY(out.alloc(sizeof(double) * (size_t)(std::min(chunk, num_snp) * num_id)));
    for (long long s0 = 0; s0 < num_snp; s0 += chunk) {
        const long long ms = std::min(chunk, num_snp - s0);
        GB_TRY(gmat_b200_decode(st.bed.as<unsigned char>() + s0 * st.row_bytes, num_id, ms, st.row_bytes,
                                nullptr, out.as<double>(), nullptr));
        GB_CUDA(cudaMemcpy(marker_mat + s0 * num_id, out.p, sizeof(double) * (size_t)(ms * num_id),
                           cudaMemcpyDeviceToHost));
    }
    return GMAT_B200_OK;
}

extern "C" int remma_epiAA_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                                   long long len_snp_lst_0, double *pymat, double eff_cut, char *out_file) {
    const Pass ps[1] = {{GMAT_B200_CODING_ADD, GMAT_B200_CODING_ADD, 0, nullptr, nullptr}};
    return run_screen(bed_file, num_id, num_snp, snp_lst_0, len_snp_lst_0, pymat, eff_cut, nullptr, ps, 1, out_file);
}

extern "C" int remma_epiDD_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                                   long long len_snp_lst_0, double *pymat, double eff_cut, char *out_file) {
    const Pass ps[1] = {{GMAT_B200_CODING_DOM, GMAT_B200_CODING_DOM, 0, nullptr, nullptr}};
    return run_screen(bed_file, num_id, num_snp, snp_lst_0, len_snp_lst_0, pymat, eff_cut, nullptr, ps, 1, out_file);
}

extern "C" int remma_epiAD_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                                   long long len_snp_lst_0, double *pymat, double eff_cut, char *out_file) {
    // direction 1: a_i * d_j, '>=' (:233-240); direction 2: d_i * a_j reported as (j, i), '>' (:241-247)
    const Pass ps[2] = {{GMAT_B200_CODING_ADD, GMAT_B200_CODING_DOM, GMAT_B200_EFF_GE, nullptr, nullptr},
                        {GMAT_B200_CODING_DOM, GMAT_B200_CODING_ADD, GMAT_B200_EFF_SWAP, nullptr, nullptr}};
    return run_screen(bed_file, num_id, num_snp, snp_lst_0, len_snp_lst_0, pymat, eff_cut, nullptr, ps, 2, out_file);
}

extern "C" int remma_epiAA_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp,
                                       long long *snp_lst_0, long long len_snp_lst_0, double *pymat,
                                       long long *freq, double *eff_cut, char *out_file) {
    GB_CHECK_ARG(freq && eff_cut, "NULL pointer");
    const Pass ps[1] = {{GMAT_B200_CODING_ADD, GMAT_B200_CODING_ADD, 0, freq, freq}};
    return run_screen(bed_file, num_id, num_snp, snp_lst_0, len_snp_lst_0, pymat, 0.0, eff_cut, ps, 1, out_file);
}

extern "C" int remma_epiDD_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp,
                                       long long *snp_lst_0, long long len_snp_lst_0, double *pymat,
                                       long long *freq, double *eff_cut, char *out_file) {
    GB_CHECK_ARG(freq && eff_cut, "NULL pointer");
    const Pass ps[1] = {{GMAT_B200_CODING_DOM, GMAT_B200_CODING_DOM, 0, freq, freq}};
    return run_screen(bed_file, num_id, num_snp, snp_lst_0, len_snp_lst_0, pymat, 0.0, eff_cut, ps, 1, out_file);
}

extern "C" int remma_epiAD_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp,
                                       long long *snp_lst_0, long long len_snp_lst_0, double *pymat,
                                       long long *freqA, long long *freqD, double *eff_cut, char *out_file) {
    GB_CHECK_ARG(freqA && freqD && eff_cut, "NULL pointer");
    // both directions index the table by freqA[loop SNP]*10 + freqD[other SNP] (:332,339)
    const Pass ps[2] = {{GMAT_B200_CODING_ADD, GMAT_B200_CODING_DOM, GMAT_B200_EFF_GE, freqA, freqD},
                        {GMAT_B200_CODING_DOM, GMAT_B200_CODING_ADD, GMAT_B200_EFF_SWAP, freqA, freqD}};
    return run_screen(bed_file, num_id, num_snp, snp_lst_0, len_snp_lst_0, pymat, 0.0, eff_cut, ps, 2, out_file);
}
