from .gmatrix import agmat, dgmat_as, output_mat
