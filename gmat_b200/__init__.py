"""gmat_b200 -- B200-native engine for GMAT's REMMAX pair scan and GRM build (drop-in for that path).

    from gmat_b200.gmatrix import agmat, dgmat_as
    from gmat_b200.remma.remma_epiAA import remma_epiAA, remma_epiAA_approx, ...

Importing the package does not touch CUDA; the first call does, and fails loudly without a GPU or
without the compiled library (there is no CPU fallback).
"""
__version__ = "0.1.0"
