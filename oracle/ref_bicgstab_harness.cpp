// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Eigen's OWN BiCGSTAB (/root/reference/lib/eigen/Eigen/src/IterativeLinearSolvers/BiCGSTAB.h:32-119, the class at :166-212,
// vendored Eigen 3.4.90) instantiated as BiCGSTAB<SparseMatrix<double>, DiagonalPreconditioner<double>> -- the solver
// SURVEY.md 8(d) names as the checker of the non-symmetric mode.  The reference application never instantiates it
// (src/FEM/CPUSolver.cpp:65 uses ConjugateGradient for every equation), so this harness is what pins
// oracle/fea_oracle_impl.h FN(bicgstab) and the CUDA BiCGSTAB kernels.  Built by oracle/Makefile into
// oracle/_ref/libeigen_bicgstab.so; only the header tree under /root/reference/lib/eigen is compiled, nothing is copied.
//
// One quirk of the vendored snapshot is handled HERE, not hidden: its loop tests `r_norm > tol` (BiCGSTAB.h:62), i.e. an
// ABSOLUTE residual, although the class documents a relative tolerance (:146).  The harness passes tol_rel * |b|
// (stableNorm, like :44) to setTolerance so that the stopping rule is the documented relative one.
#include <cstdint>

#include <Eigen/IterativeLinearSolvers>
#include <Eigen/Sparse>

extern "C" {

// A in CSR (n rows, rowptr / col / val), b, x out.  Returns Eigen's info() (0 = Success).  iters / err = iterations(), error().
int eigen_bicgstab_solve(int64_t n, const int32_t* rowptr, const int32_t* col, const double* val, const double* b, double tol_rel,
                         int64_t maxit, double* x, int64_t* iters, double* err) {
    using RowMat = Eigen::SparseMatrix<double, Eigen::RowMajor, int>;
    Eigen::Map<const RowMat> Ar((Eigen::Index)n, (Eigen::Index)n, (Eigen::Index)rowptr[n], rowptr, col, val);
    Eigen::SparseMatrix<double> A = Ar;  // column-major, as SparseMatrix<double> is declared in the survey's checker
    Eigen::Map<const Eigen::VectorXd> rhs(b, (Eigen::Index)n);
    Eigen::BiCGSTAB<Eigen::SparseMatrix<double>, Eigen::DiagonalPreconditioner<double>> solver(A);
    solver.setTolerance(tol_rel * rhs.stableNorm());
    if (maxit >= 0) solver.setMaxIterations((Eigen::Index)maxit);
    Eigen::VectorXd sol = solver.solve(rhs);  // x0 = 0
    for (int64_t i = 0; i < n; i++) x[i] = sol[i];
    if (iters) *iters = (int64_t)solver.iterations();
    if (err) *err = (double)solver.error();
    return (int)solver.info();
}
}
