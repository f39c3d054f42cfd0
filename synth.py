"""Synthetic PLINK data sets of a named n x m shape for tests and bench.py.

Genotypes: per-SNP allele frequency p ~ U(0.05, 0.95), g ~ Binomial(2, p) i.i.d. (HWE, no LD), no missing
calls unless ``missing_rate`` > 0; written as a real PLINK1 SNP-major .bed (+ .bim/.fam/pheno) so that the CUDA
engine and the CPU reference read identical bytes.  Phenotype: y = 1 + covariate + X_A beta + planted
epistatic pairs + e.
"""
import os

import numpy as np

CODE_OF_GENO = np.array([0, 2, 3], dtype=np.uint8)      # g=0 -> 00, g=1 -> 10, g=2 -> 11; missing -> 01


def pack_bed(geno_snp_major):
    """[m][n] int array of 0/1/2 (or -1 = missing) -> PLINK payload bytes [m][ceil(n/4)]."""
    g = np.asarray(geno_snp_major)
    m, n = g.shape
    codes = np.where(g < 0, 1, CODE_OF_GENO[np.clip(g, 0, 2)]).astype(np.uint8)
    pad = (-n) % 4
    if pad:
        codes = np.concatenate([codes, np.zeros((m, pad), dtype=np.uint8)], axis=1)
    c = codes.reshape(m, -1, 4)
    return (c[:, :, 0] | (c[:, :, 1] << 2) | (c[:, :, 2] << 4) | (c[:, :, 3] << 6)).astype(np.uint8)


def random_genotypes(n, m, seed, missing_rate=0.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    p = rng.uniform(0.05, 0.95, size=m)
    g = rng.binomial(2, p[:, None], size=(m, n)).astype(np.int8)
    for s in range(m if n >= 2 else 0):                 # no monomorphic columns (impossible for n == 1)
        for _ in range(1000):
            if g[s].min() != g[s].max():
                break
            g[s] = rng.binomial(2, p[s], size=n)
        else:                                           # tiny n with an extreme p: force one heterozygote
            g[s, 0] = 1 if g[s, 0] != 1 else 0
    if missing_rate > 0:
        g[rng.uniform(size=(m, n)) < missing_rate] = -1
    return g


def write_dataset(prefix, n, m, seed=20201013, missing_rate=0.0, records_per_id=1, n_planted=10,
                  planted_kind="AA", geno=None, also_records=()):
    """Write prefix.bed/.bim/.fam and prefix.pheno; returns dict(geno=[m][n] int8, planted=[(i,j)], pheno=path).
    planted_kind "mix" cycles AA, AD, DD over the planted pairs; also_records=(R, ...) writes prefix.rep<R>.pheno
    with R records per individual (same genetic values plus a permanent-environment effect) beside prefix.pheno."""
    rng = np.random.Generator(np.random.PCG64(seed + 1))
    g = random_genotypes(n, m, seed, missing_rate) if geno is None else np.asarray(geno, dtype=np.int8)
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    with open(prefix + ".bed", "wb") as f:
        f.write(bytes([0x6c, 0x1b, 0x01]))
        f.write(pack_bed(g).tobytes())
    with open(prefix + ".bim", "w") as f:
        for j in range(m):
            f.write("1 snp%d 0 %d A T\n" % (j, j + 1))
    with open(prefix + ".fam", "w") as f:
        for k in range(n):
            f.write("fam%d id%d 0 0 0 -9\n" % (k, k))
    gz = np.where(g < 0, 0, g).astype(np.float64)
    freq = gz.sum(axis=1, keepdims=True) / (2 * n)
    xa = gz - 2 * freq
    h = np.where(gz > 1.5, 0.0, gz)
    xd = h - 2 * freq * (1 - freq)
    sel = rng.choice(m, size=min(100, m), replace=False)
    gv = (xa[sel] * rng.normal(0, 0.05, size=(len(sel), 1))).sum(axis=0)
    planted = []
    for q in range(n_planted if m >= 4 else 0):
        i, j = sorted(rng.choice(m, size=2, replace=False).tolist())
        kind_q = ("AA", "AD", "DD")[q % 3] if planted_kind == "mix" else planted_kind
        u = xa[i] if kind_q in ("AA", "AD") else xd[i]
        v = xa[j] if kind_q == "AA" else xd[j]
        w = u * v
        gv = gv + w * (rng.uniform(0.25, 0.4) / max(w.std(), 1e-9)) * rng.choice([-1.0, 1.0])
        planted.append((int(i), int(j)))
    cov = rng.integers(0, 2, size=n)
    perm = rng.normal(0, np.sqrt(0.3), size=n) if records_per_id > 1 else np.zeros(n)
    with open(prefix + ".pheno", "w") as f:
        for k in range(n):
            for _ in range(records_per_id):
                yv = 1.0 + 0.3 * cov[k] + gv[k] + perm[k] + rng.normal()
                f.write("fam%d id%d 1 %d %.6f\n" % (k, k, cov[k], yv))
    out = {"geno": g, "planted": planted, "pheno": prefix + ".pheno"}
    for reps in also_records:
        pe = rng.normal(0, np.sqrt(0.3), size=n)
        path = prefix + ".rep%d.pheno" % reps
        with open(path, "w") as f:
            for k in range(n):
                for _ in range(reps):
                    f.write("fam%d id%d 1 %d %.6f\n" % (k, k, cov[k], 1.0 + 0.3 * cov[k] + gv[k] + pe[k] + rng.normal()))
        out["pheno_rep%d" % reps] = path
    return out


def device_payload(n, m, seed, device, row_bytes, chunk=8192):
    """Random genotype store generated directly in HBM (for shapes too large to make on the host, e.g.
    config C2 n=10,000 x m=500,000): uint8 [m][row_bytes] tensor, HWE genotypes, no missing."""
    import torch
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    out = torch.zeros((m, row_bytes), dtype=torch.uint8, device=device)
    npad = (n + 3) // 4 * 4
    for s0 in range(0, m, chunk):
        ms = min(chunk, m - s0)
        p = torch.rand((ms, 1), generator=gen, device=device) * 0.9 + 0.05
        u = torch.rand((ms, npad), generator=gen, device=device)
        g = (u < p * p).to(torch.uint8) * 3                                    # 11 = two copies
        g = torch.where((u >= p * p) & (u < p * p + 2 * p * (1 - p)), torch.full_like(g, 2), g)   # 10 = het
        if npad != n:
            g[:, n:] = 0
        g = g.view(ms, npad // 4, 4)
        out[s0:s0 + ms, :npad // 4] = g[:, :, 0] | (g[:, :, 1] << 2) | (g[:, :, 2] << 4) | (g[:, :, 3] << 6)
    return out
