"""Drop-in for gmat/uvlmm/uvlmm_varcom.py:8-126: REML variance components of the
univariate mixed model with several relationship matrices, weighted EM + AI updates -- the producer of the
``var_com`` every remma_epiXX call takes.

Same signatures, logging and output file.  The arithmetic runs in float64 on the GPU (torch: cuSOLVER for the
factorisation / inverse of V, cuSPARSE for the products with the incidence matrix Z); per iteration the
reference spends 2 N^3 (slogdet + inv) + r N^3 (trace(P ZG_kZ') formed through a full product,
uvlmm_varcom.py:66); here V is factorised once (Cholesky: log-determinant and inverse from the same factor)
and the traces are taken as sum(P o ZG_kZ') -- N^2 each.  Convergence control (EM weight search, criteria)
is the reference's, on the host.  There is no CPU fallback.
"""
import logging

import numpy as np
import torch

from .. import engine
from .design_matrix import design_matrix_wemai_multi_gmat


def _wemai_multi_gmat(y, xmat, zmat, gmat_lst, init=None, maxiter=200, cc_par=1.0e-8, cc_gra=1.0e-6):
    device = engine.current_device()
    f64 = torch.float64
    logging.info("#####Prepare#####")
    var_com = [1.0] * (len(gmat_lst) + 1)
    if init is not None:
        var_com = init[:]
    var_com = np.array(var_com, dtype=np.float64)
    var_com_new = var_com[:]
    yv = torch.as_tensor(np.ascontiguousarray(np.array(y, dtype=np.float64).reshape(-1, 1))).to(device)
    n = yv.shape[0]
    xm = torch.as_tensor(np.ascontiguousarray(np.array(xmat, dtype=np.float64).reshape(n, -1))).to(device)
    zs, _ = engine.sparse_incidence(zmat, device)
    zgmat_lst = []
    for k in range(len(gmat_lst)):
        g = torch.as_tensor(np.ascontiguousarray(gmat_lst[k], dtype=np.float64)).to(device)
        zgmat_lst.append(torch.sparse.mm(zs, torch.sparse.mm(zs, g).t().contiguous()))      # Z (Z G)'  (:34)
        del g
    logging.info('Initial variances: ' + ' '.join(list(np.array(var_com_new, dtype=str))))
    logging.info("#####Start the iteration#####\n\n")
    it = 0
    cc_gra_val = 1000.0
    cc_par_val = 1000.0
    eye = torch.eye(n, dtype=f64, device=device)
    while it < maxiter:
        it += 1
        logging.info('###Round: ' + str(it) + '###')
        vmat = eye * float(var_com[-1])
        for k in range(len(zgmat_lst)):
            vmat += zgmat_lst[k] * float(var_com[k])
        chol, info = torch.linalg.cholesky_ex(vmat)
        if int(info.item()) == 0:
            ll_v = 2.0 * float(torch.log(torch.diagonal(chol)).sum().item())
            vinv = torch.cholesky_inverse(chol)
        else:                                       # V not positive definite: the reference's LU route
            ll_v = float(torch.linalg.slogdet(vmat)[1].item())
            vinv = torch.linalg.inv(vmat)
        del vmat, chol
        vx = vinv @ xm
        xvx = xm.t() @ vx
        ll_xvx = float(torch.linalg.slogdet(xvx)[1].item())
        xvx = torch.linalg.inv(xvx)
        pmat = vinv - vx @ xvx @ vx.t()
        del vinv
        pymat = pmat @ yv
        ll_ypy = float((yv.t() @ pymat).sum().item())
        ll_val = -2 * (ll_v + ll_xvx + ll_ypy)
        fd_mat, wv = [], []
        for k in range(len(zgmat_lst)):
            zgpy = zgmat_lst[k] @ pymat
            # trace(P ZGZ) = sum(P o ZGZ') (ZGZ is symmetric up to rounding; the transpose keeps the identity exact)
            fd = -float((pmat * zgmat_lst[k].t()).sum().item()) + float((pymat.t() @ zgpy).sum().item())
            fd_mat.append(0.5 * fd)
            wv.append(zgpy)
        fd = -float(torch.diagonal(pmat).sum().item()) + float((pymat.t() @ pymat).sum().item())
        fd_mat.append(0.5 * fd)
        fd_mat = np.array(fd_mat)
        wv.append(pymat)
        wv = torch.cat(wv, dim=1)
        ai_mat = (0.5 * (wv.t() @ pmat @ wv)).cpu().numpy()
        del pymat, wv, pmat
        em_mat = np.diag([n / (v * v) for v in var_com])
        delta = None
        for j in range(0, 101):
            weight = j * 0.01
            wemai_mat = (1 - weight) * ai_mat + weight * em_mat
            delta = np.dot(np.linalg.inv(wemai_mat), fd_mat)
            var_com_new = var_com + delta
            if min(var_com_new) > 0:
                logging.info('EM weight value: ' + str(weight))
                break
        cc_par_val = np.sqrt(np.sum(delta * delta) / np.sum(var_com_new * var_com_new))
        var_com = var_com_new[:]
        cc_gra_val = np.sqrt(np.sum(fd_mat * fd_mat))
        logging.info('-2logL for last iteration: ' + str(ll_val))
        logging.info('Norm of gradient vector for last iteration: ' + str(cc_gra_val))
        logging.info('Norm of update vector: ' + str(cc_par_val))
        logging.info('Updated variances: ' + ' '.join(list(np.array(var_com_new, dtype=str))))
        if cc_gra_val < cc_gra and cc_par_val < cc_par:
            break
    if cc_gra_val < cc_gra and cc_par_val < cc_par:
        logging.info('Variances converged.')
    else:
        logging.info('Variances not converged.')
    return var_com


def wemai_multi_gmat(pheno_file, bed_file, gmat_lst, init=None, maxiter=200, cc_par=1.0e-8, cc_gra=1.0e-6,
                     out_file='wemai_multi_gmat.var'):
    """uvlmm_varcom.py:107-126: variances from files; writes them to ``out_file`` with np.savetxt."""
    y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
    var_com = _wemai_multi_gmat(y, xmat, zmat, gmat_lst, init=init, maxiter=maxiter, cc_par=cc_par, cc_gra=cc_gra)
    np.savetxt(out_file, var_com)
    return var_com
