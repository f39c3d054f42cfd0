"""Builds the reference-named front-ends (remma_epiXX, _parallel, _pair, _eff, _eff_parallel, _maf_eff,
_approx, _approx_parallel, _maf_approx, ... and the underscore forms taking (y, xmat, zmat)) for one kind.
Signatures, defaults and return value (0) follow gmat/remma/remma_epi{AA,AD,DD}/*.py."""
from ..uvlmm.design_matrix import design_matrix_wemai_multi_gmat
from . import _core


def build(kind, ns):
    K = kind.name
    dflt = {'AA': 1.0e-5, 'AD': 0.0001, 'DD': 0.0001}[K]       # underscore-form p_cut defaults differ per copy

    def _scan(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, p_cut=dflt, out_file='epi' + K):
        return _core.exact_scan(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0, p_cut, out_file)

    def scan(pheno_file, bed_file, gmat_lst, var_com, snp_lst_0=None, p_cut=1.0e-5, out_file='epi' + K):
        import time
        t0 = time.perf_counter()
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
        _core.PHASES["design_matrix"] = time.perf_counter() - t0
        return _scan(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_lst_0, p_cut=p_cut, out_file=out_file)

    def _scan_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, p_cut=1.0e-5 if K == 'AA' else 1.0e-4,
                       out_file='epi%s_parallel' % K):
        return _core.exact_scan_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, p_cut, out_file)

    def scan_parallel(pheno_file, bed_file, gmat_lst, var_com, parallel, p_cut=1.0e-5, out_file='epi%s_parallel' % K):
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
        return _scan_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, p_cut=p_cut, out_file=out_file)

    def _pair(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_pair_file, max_test_pair=50000, p_cut=1.0e-4,
              out_file='epi%s_pair' % K):
        return _core.pair_test(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_pair_file, max_test_pair,
                               p_cut, out_file)

    def pair(pheno_file, bed_file, gmat_lst, var_com, snp_pair_file, max_test_pair=50000, p_cut=1.0e-4,
             out_file='epi%s_pair' % K):
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
        return _pair(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_pair_file, max_test_pair=max_test_pair,
                     p_cut=p_cut, out_file=out_file)

    def _eff(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, var_app=1.0, p_cut=1.0e-5,
             out_file='epi%s_eff' % K):
        return _core.eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0, var_app, p_cut, out_file)

    def eff(pheno_file, bed_file, gmat_lst, var_com, snp_lst_0=None, var_app=1.0, p_cut=1.0e-5,
            out_file='epi%s_eff' % K):
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
        return _eff(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_lst_0, var_app=var_app, p_cut=p_cut,
                    out_file=out_file)

    def _eff_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, var_app=1.0, p_cut=1.0e-5,
                      out_file='epi%s_eff_parallel' % K):
        return _core.eff_screen_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, var_app,
                                         p_cut, out_file)

    def eff_parallel(pheno_file, bed_file, gmat_lst, var_com, parallel, var_app=1.0, p_cut=1.0e-5,
                     out_file='epi%s_eff_parallel' % K):
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
        return _eff_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, var_app=var_app, p_cut=p_cut,
                             out_file=out_file)

    if K == 'AD':
        def _maf_eff(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, freqA=None, freqD=None,
                     freq_deno=None, p_cut=1.0e-5, out_file='epiAD_maf_eff'):
            return _core.maf_eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0, freqA, freqD,
                                        freq_deno, p_cut, out_file)

        def maf_eff(pheno_file, bed_file, gmat_lst, var_com, snp_lst_0=None, freqA=None, freqD=None, freq_deno=None,
                    p_cut=1.0e-5, out_file='epiAA_maf_eff'):
            y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
            return _maf_eff(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_lst_0, freqA=freqA,
                            freqD=freqD, freq_deno=freq_deno, p_cut=p_cut, out_file=out_file)
    else:
        # (the AA underscore form keeps the _eff default name: remma_epiAA_maf_eff.py:20)
        def _maf_eff(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, freq=None, freq_deno=None,
                     p_cut=1.0e-5, out_file='epiAA_eff' if K == 'AA' else 'epi%s_maf_eff' % K):
            return _core.maf_eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0, freq, None,
                                        freq_deno, p_cut, out_file)

        def maf_eff(pheno_file, bed_file, gmat_lst, var_com, snp_lst_0=None, freq=None, freq_deno=None,
                    p_cut=1.0e-5, out_file='epi%s_maf_eff' % K):
            y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
            return _maf_eff(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_lst_0, freq=freq,
                            freq_deno=freq_deno, p_cut=p_cut, out_file=out_file)

    if K == 'AD':
        # (the AD copy keeps the AA default name: remma_epiAD_maf_eff.py:131-132,167-168)
        def _maf_eff_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freqA=None, freqD=None,
                              freq_deno=None, p_cut=1.0e-5, out_file='epiAA_maf_eff_parallel'):
            return _core.maf_eff_screen_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freqA,
                                                 freqD, freq_deno, p_cut, out_file)

        def maf_eff_parallel(pheno_file, bed_file, gmat_lst, var_com, parallel, freqA=None, freqD=None,
                             freq_deno=None, p_cut=1.0e-5, out_file='epiAA_maf_eff_parallel'):
            y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
            return _maf_eff_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freqA=freqA, freqD=freqD,
                                     freq_deno=freq_deno, p_cut=p_cut, out_file=out_file)
    else:
        def _maf_eff_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freq=None, freq_deno=None,
                              p_cut=1.0e-5, out_file='epi%s_maf_eff_parallel' % K):
            return _core.maf_eff_screen_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freq,
                                                 None, freq_deno, p_cut, out_file)

        def maf_eff_parallel(pheno_file, bed_file, gmat_lst, var_com, parallel, freq=None, freq_deno=None,
                             p_cut=1.0e-5, out_file='epi%s_maf_eff_parallel' % K):
            y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
            return _maf_eff_parallel(y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freq=freq,
                                     freq_deno=freq_deno, p_cut=p_cut, out_file=out_file)

    def approx(pheno_file, bed_file, gmat_lst, var_com, p_cut=1.0e-5, num_random_pair=100000,
               out_file='epi%s_approx' % K):
        return _core.approx(kind, pheno_file, bed_file, gmat_lst, var_com, p_cut, num_random_pair, out_file)

    def approx_parallel(pheno_file, bed_file, gmat_lst, var_com, parallel, p_cut=1.0e-5, num_random_pair=100000,
                        out_file='epi%s_approx' % K):
        return _core.approx(kind, pheno_file, bed_file, gmat_lst, var_com, p_cut, num_random_pair, out_file,
                            parallel=parallel)

    def maf_approx(pheno_file, bed_file, gmat_lst, var_com, p_cut=1.0e-5, num_random_pair=100000,
                   out_file='epi%s_maf_approx' % K):
        return _core.approx(kind, pheno_file, bed_file, gmat_lst, var_com, p_cut, num_random_pair, out_file, maf=True)

    def maf_approx_parallel(pheno_file, bed_file, gmat_lst, var_com, parallel, p_cut=1.0e-5, num_random_pair=100000,
                            out_file='epi%s_maf_approx_parallel' % K):
        return _core.approx(kind, pheno_file, bed_file, gmat_lst, var_com, p_cut, num_random_pair, out_file,
                            maf=True, parallel=parallel)

    table = {
        'remma_epi%s': scan, '_remma_epi%s': _scan,
        'remma_epi%s_parallel': scan_parallel, '_remma_epi%s_parallel': _scan_parallel,
        'remma_epi%s_pair': pair, '_remma_epi%s_pair': _pair,
        'remma_epi%s_eff': eff, '_remma_epi%s_eff': _eff,
        'remma_epi%s_eff_parallel': eff_parallel, '_remma_epi%s_eff_parallel': _eff_parallel,
        'remma_epi%s_maf_eff': maf_eff, '_remma_epi%s_maf_eff': _maf_eff,
        'remma_epi%s_maf_eff_parallel': maf_eff_parallel, '_remma_epi%s_maf_eff_parallel': _maf_eff_parallel,
        'remma_epi%s_approx': approx, 'remma_epi%s_approx_parallel': approx_parallel,
        'remma_epi%s_maf_approx': maf_approx, 'remma_epi%s_maf_approx_parallel': maf_approx_parallel,
    }
    for pat, fn in table.items():
        name = pat % K
        fn.__name__ = name
        fn.__qualname__ = name
        ns[name] = fn
    return sorted(pat % K for pat in table)
