"""Drop-in for ``annotation_snp_pos`` (gmat/remma/annotation.py:22-74): step 4 of the README workflow
(README.md:101-119) -- join the rows of a remma_epiXX result file with the .bim records of both SNPs, keep
rows with p <= p_cut whose SNPs lie on different chromosomes or more than ``dis`` bp apart, optionally drop
pairs in LD.  Pure host text glue around the hot path; same output files (``res_file.anno``, ``.anno.ld``).
"""

_COLS = ('chr', 'ID', 'cm', 'bp', 'allele1', 'allele2')


def annotation_snp_pos(res_file, bed_file, p_cut=1, dis=0, ld_file=None, r2=0.2):
    with open(bed_file + '.bim') as fin:
        bim = [line.split() for line in fin]                  # row k = the SNP the result files call "k"
    anno_file = res_file + '.anno'
    with open(res_file) as fin, open(anno_file, 'w') as fout:
        head = fin.readline().split()
        fout.write(' '.join([head[0]] + ['snp0_' + c for c in _COLS] + [head[1]] + ['snp1_' + c for c in _COLS]
                            + head[2:]) + '\n')
        for line in fin:
            arr = line.split()
            s0, s1 = bim[int(arr[0])], bim[int(arr[1])]
            apart = s0[0] != s1[0] or abs(float(s0[3]) - float(s1[3])) > dis
            if float(arr[-1]) <= p_cut and apart:
                fout.write(' '.join([arr[0]] + s0 + [arr[1]] + s1 + arr[2:]) + '\n')
    if ld_file is not None:
        linked = set()
        with open(ld_file) as fin:
            fin.readline()
            for line in fin:
                arr = line.split()
                if float(arr[-1]) > r2:                       # plink --r2 table: SNP_A = col 2, SNP_B = col 5
                    linked.add((arr[2], arr[5]))
                    linked.add((arr[5], arr[2]))
        with open(anno_file) as fin, open(anno_file + '.ld', 'w') as fout:
            fout.write(fin.readline())
            for line in fin:
                arr = line.split()
                if (arr[2], arr[9]) not in linked:
                    fout.write(line)
    return 0
