// cfb200 — SURVEY.md §8(f) rank 3: consumers of the device-resident block Jacobian that never take it through the host.
//   * BsrToCsr: the scalar CSR Paralution wants (ParalutionMatrix.hh:144 SetDataPtrCSR, ParalutionMatrix.cu:48-105) and PETSc
//     MatCreateSeqAIJWithArrays-style consumers, written on the device (replaces the host loop + 14.8 GB download of
//     cfb200_export_csr);
//   * a restarted GMRES with a block-Jacobi preconditioner (the diagonal q x q blocks inverted once per solve), the role
//     PETSc KSPGMRES + PCBJACOBI play behind LinearSystemSolver::solveSys (plugins/Petsc/StdParSolveSys.cxx): J x = rhs
//     with x returned in the rhs buffer, which is where NewtonMethod's StdUpdateSol reads dU (StdUpdateSol.cxx:110-141).
// Everything is deterministic: row sums run in storage order, dot products use the fixed two-level block sums of
// CudaBackend::sumOrdered, no atomics.  Included by cfb_capi.cu only (CUDA build).
#pragma once

namespace cfb {

// ---- BSR -> scalar CSR: one thread per block entry (k, ib, jb) -------------------------------------------------------
// rows of a cell one after the other, each with its columns ascending (the layout of cfb200_export_csr)
template <typename IDX>
struct BsrToCsr {   // idx over one chunk of nnzb*q*q entries
  const int* rowPtr; const int* colInd; const int* blockRow; const double* bsr; int q; long long base;
  IDX* rowOff; IDX* col; double* val;
  __device__ void operator()(int idx0) const {
    const long long idx = base + idx0;
    const int qq = q*q;
    const long long k = idx/qq;
    const int e = (int)(idx - k*qq), ib = e/q, jb = e - ib*q;
    const int row = blockRow[k];
    const long long first = rowPtr[row], nb = rowPtr[row+1] - first;
    // scalar row (row*q + ib) starts after the ib earlier scalar rows of this block row and all earlier block rows
    const long long p = first*qq + (long long)ib*nb*q + (k - first)*q + jb;
    if (val) val[p] = bsr[idx];
    if (col) col[p] = (IDX)((long long)colInd[k]*q + jb);
    if (rowOff && jb == 0 && k == first) rowOff[(long long)row*q + ib] = (IDX)(first*qq + (long long)ib*nb*q);
  }
};

// ---- GMRES building blocks ---------------------------------------------------------------------------------------------
// y = A x, one thread per scalar row; the row sum runs over the blocks in storage order (columns ascending) and over jb
struct BsrSpmv {
  const int* rowPtr; const int* colInd; const double* bsr; const double* x; double* y; int q;
  __device__ void operator()(int r) const {
    const int row = r/q, ib = r - row*q;
    double acc = 0.;
    for (int k = rowPtr[row]; k < rowPtr[row+1]; ++k) {
      const double* blk = bsr + ((size_t)k*q + ib)*q;
      const double* xs = x + (size_t)colInd[k]*q;
      for (int jb = 0; jb < q; ++jb) acc += blk[jb]*xs[jb];
    }
    y[r] = acc;
  }
};
// inverse of every diagonal block: Gauss-Jordan without pivoting (the pseudo-time term makes the diagonal blocks strongly
// diagonally dominant); rows of non-updatable cells get the identity.  One thread per cell, q <= 9.
struct InvertDiag {      // idx over block rows (the matrix's own numbering)
  const int* rowPtr; const int* colInd; const uint8_t* updatable; const double* bsr; double* inv; int q;
  __device__ void operator()(int c) const {
    double a[81], b[81];
    const int qq = q*q;
    int ds = rowPtr[c];
    while (colInd[ds] != c) ++ds;                       // columns ascending, the diagonal block is always present
    const double* blk = bsr + (size_t)ds*qq;
    for (int i = 0; i < qq; ++i) { a[i] = blk[i]; b[i] = 0.; }
    for (int i = 0; i < q; ++i) b[i*q + i] = 1.;
    if (updatable[c]) {
      for (int p = 0; p < q; ++p) {
        const double ip = 1.0/a[p*q + p];
        for (int j = 0; j < q; ++j) { a[p*q + j] = a[p*q + j]*ip; b[p*q + j] = b[p*q + j]*ip; }
        for (int i = 0; i < q; ++i) {
          if (i == p) continue;
          const double f = a[i*q + p];
          for (int j = 0; j < q; ++j) { a[i*q + j] = a[i*q + j] - f*a[p*q + j]; b[i*q + j] = b[i*q + j] - f*b[p*q + j]; }
        }
      }
    }
    double* o = inv + (size_t)c*qq;
    for (int i = 0; i < qq; ++i) o[i] = b[i];
  }
};
// y = M^-1 x with the inverted diagonal blocks (or a copy when there is no preconditioner)
struct ApplyDiagInv {
  const double* inv; const double* x; double* y; int q; int on;
  __device__ void operator()(int r) const {
    if (!on) { y[r] = x[r]; return; }
    const int row = r/q, ib = r - row*q;
    const double* blk = inv + ((size_t)row*q + ib)*q;
    const double* xs = x + (size_t)row*q;
    double acc = 0.;
    for (int jb = 0; jb < q; ++jb) acc += blk[jb]*xs[jb];
    y[r] = acc;
  }
};
struct VecMul { const double* a; const double* b; double* o; __device__ void operator()(int i) const { o[i] = a[i]*b[i]; } };
struct VecAxpy { double alpha; const double* x; double* y; __device__ void operator()(int i) const { y[i] = y[i] + alpha*x[i]; } };
struct VecScaleTo { double alpha; const double* x; double* y; __device__ void operator()(int i) const { y[i] = alpha*x[i]; } };
struct VecZeroMasked {   // rows of non-updatable cells carry no equation: their right-hand side is dropped
  const uint8_t* updatable; double* x; int q;
  __device__ void operator()(int i) const { if (!updatable[i/q]) x[i] = 0.; }
};

} // namespace cfb
