/* TEST INFRASTRUCTURE ONLY -- the oracle.  Included twice by fea_oracle.c, once with REAL = float
 * (SUF = 32) and once with REAL = double (SUF = 64).
 *
 * A plain-C restatement, operation by operation, of the reference's CPU hot path
 *   src/FEM/FEMContext.cpp:70-240   (idx_map, P1 stiffness / mass / advection assembly)
 *   src/FEM/CPUSolver.cpp:13-147    (gather/scatter, per-equation implicit-Euler step)
 *   lib/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:30-85  (Jacobi-PCG recurrence)
 *   lib/eigen/Eigen/src/IterativeLinearSolvers/BasicPreconditioners.h:63-86 (DiagonalPreconditioner)
 *   lib/eigen/Eigen/src/IterativeLinearSolvers/BiCGSTAB.h:31-119 (the "correct" non-symmetric mode)
 *   lib/eigen/Eigen/src/SparseCore/SparseMatrix.h:1125-1173, 1475-1511 (setFromTriplets semantics)
 * (paths relative to /root/reference).  Vertex coordinates and parameters are the reference's float
 * values promoted to REAL; every arithmetic step is then done in REAL in the reference's order.
 * The fp32 instantiation is pinned against the compiled reference (oracle/_ref, tests/test_oracle_pin.py);
 * the fp64 instantiation is the same code with REAL = double.
 */

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(CAT(orc, SUF), CAT(_, name))
#define HANDLE CAT(OrcHandle, SUF)

typedef struct {
    int64_t n, nnz;
    int32_t* outer; /* n + 1 */
    int32_t* inner; /* nnz, sorted ascending within each outer vector */
    REAL* val;
} CAT(Csc, SUF);
#define CSC CAT(Csc, SUF)

typedef struct {
    int64_t V, T;
    float* xyz;          /* V x 3, the reference's glm::vec3 storage */
    uint32_t* tri;       /* T x 3 */
    uint8_t* on_boundary;
    int64_t num_boundary;
    int bc, equation;
    int32_t* idx_map;    /* V */
    int64_t n;           /* num_unknowns */
    CSC M, K, Adv;       /* Adv may have a smaller pattern than M/K (Neumann + open mesh) */
    REAL* adv_on_m;      /* Adv values scattered on M's pattern, 0 where absent */
    float par[4][5];     /* per equation {dt, c0..c3} in the reference's float type */
    float velocity[3];
    float* values;       /* V, the reference's surface->values (always float) */
    REAL *u, *v;         /* n, solver state (CPUSolver.hpp:17-18) */
    int solver_mode;     /* 0 = parity (PCG multiplying by A^T, as Eigen does), 1 = BiCGSTAB for adv-diff */
    double tol;          /* <0: epsilon of REAL (IterativeSolverBase.h:368) */
    int64_t max_iter;    /* <0: 2n (IterativeSolverBase.h:251) */
    int64_t last_iters[2];
    double last_err[2];
} HANDLE;

/* ---------- Eigen 3-vector helpers: fixed-size redux of 3 terms is t0 + (t1 + t2) ----------
 * (redux_novec_unroller splits [0,3) into [0,1) and [1,3), lib/eigen/Eigen/src/Core/Redux.h) */
static inline REAL FN(sum3)(REAL a, REAL b, REAL c) { return a + (b + c); }
/* ... but dot() of fixed 3-vectors goes through inner_product_impl<Evaluator,false>
 * (lib/eigen/Eigen/src/Core/InnerProduct.h:109-121): result = a0*b0; result = pmadd(a_k, b_k, result),
 * and the SCALAR pmadd is numext::fma, a true fused multiply-add (GenericPacketMath.h:1344-1348,
 * MathFunctions.h:953-958) -- probed: bit-exact against Eigen on 1e5 random vectors. */
static inline REAL FN(fma_)(REAL a, REAL b, REAL c) { return sizeof(REAL) == 4 ? (REAL)fmaf((float)a, (float)b, (float)c) : (REAL)fma((double)a, (double)b, (double)c); }
static inline REAL FN(edot3)(const REAL* a, const REAL* b) { return FN(fma_)(a[2], b[2], FN(fma_)(a[1], b[1], a[0] * b[0])); }

/* glm::cross / glm::dot / glm::length / glm::normalize (lib/glm/glm/detail/func_geometric.inl):
 * dot = tmp.x + tmp.y + tmp.z (left to right), normalize = v * inversesqrt(dot(v,v)), inversesqrt = 1/sqrt */
static inline void FN(glm_cross)(const REAL* x, const REAL* y, REAL* o) {
    o[0] = x[1] * y[2] - y[1] * x[2];
    o[1] = x[2] * y[0] - y[2] * x[0];
    o[2] = x[0] * y[1] - y[0] * x[1];
}
static inline REAL FN(glm_dot)(const REAL* a, const REAL* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

static REAL FN(rsqrt_)(REAL x) { return (REAL)1 / (sizeof(REAL) == 4 ? (REAL)sqrtf((float)x) : (REAL)sqrt((double)x)); }
static REAL FN(sqrt_)(REAL x) { return sizeof(REAL) == 4 ? (REAL)sqrtf((float)x) : (REAL)sqrt((double)x); }

/* Per-triangle quantities shared by the three assembly loops */
typedef struct {
    REAL normal[3]; /* glm::normalize(cross) FEMContext.cpp:76 */
    REAL area;      /* 0.5f * glm::length(cross)   :77 */
    REAL jdet;      /* Eigen cross().norm()        :126, :163 */
    REAL G[3][3];   /* physical gradients           :93-95 */
} CAT(Elem, SUF);
#define ELEM CAT(Elem, SUF)

static void FN(element)(const HANDLE* h, int64_t k, ELEM* e) {
    REAL a[3], b[3], c[3], e1[3], e2[3], w[3];
    for (int d = 0; d < 3; d++) {
        a[d] = (REAL)h->xyz[3 * (int64_t)h->tri[3 * k + 0] + d];
        b[d] = (REAL)h->xyz[3 * (int64_t)h->tri[3 * k + 1] + d];
        c[d] = (REAL)h->xyz[3 * (int64_t)h->tri[3 * k + 2] + d];
        e1[d] = b[d] - a[d];
        e2[d] = c[d] - a[d];
    }
    FN(glm_cross)(e1, e2, w);
    REAL ww = FN(glm_dot)(w, w);
    REAL inv = FN(rsqrt_)(ww);
    for (int d = 0; d < 3; d++) e->normal[d] = w[d] * inv;
    e->area = (REAL)0.5 * FN(sqrt_)(ww);

    /* Eigen Vector3::cross then norm(): same component formulas (lhs1*rhs2 - lhs2*rhs1 ...),
     * squaredNorm = abs2().sum() = x^2 + (y^2 + z^2) by the unrolled redux (unlike dot(), see dot3). */
    REAL cx = e1[1] * e2[2] - e1[2] * e2[1];
    REAL cy = e1[2] * e2[0] - e1[0] * e2[2];
    REAL cz = e1[0] * e2[1] - e1[1] * e2[0];
    e->jdet = FN(sqrt_)(FN(sum3)(cx * cx, cy * cy, cz * cz));

    /* J = [e1 | e2 | normal] (FEMContext.cpp:85-89); Eigen 3x3 inverse by cofactors
     * (lib/eigen/Eigen/src/LU/InverseImpl.h:107-142): inv(i,j) = cofactor<j,i>(J) * (1/det),
     * det = sum(cofactors_col0 .* J.col(0)); then transposeInPlace, so At(i,j) = cofactor<i,j> * invdet */
    REAL m[3][3];
    for (int d = 0; d < 3; d++) {
        m[d][0] = e1[d];
        m[d][1] = e2[d];
        m[d][2] = e->normal[d];
    }
    REAL cof[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
            cof[i][j] = m[i1][j1] * m[i2][j2] - m[i1][j2] * m[i2][j1];
        }
    REAL det = FN(sum3)(cof[0][0] * m[0][0], cof[1][0] * m[1][0], cof[2][0] * m[2][0]);
    REAL invdet = (REAL)1 / det;
    REAL At[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) At[i][j] = cof[i][j] * invdet;
    /* physical_gradients[i] = At * reference_gradients[i], coefficient = t0 + (t1 + t2) */
    static const REAL rg[3][3] = {{-1, -1, 0}, {1, 0, 0}, {0, 1, 0}};
    for (int i = 0; i < 3; i++)
        for (int r = 0; r < 3; r++) e->G[i][r] = FN(sum3)(At[r][0] * rg[i][0], At[r][1] * rg[i][1], At[r][2] * rg[i][2]);
}

/* ---------- setFromTriplets: column-major, sorted inner indices, duplicates summed in triplet
 * order, explicit zeros kept (SparseMatrix.h:1125-1173 / collapseDuplicates :1475-1511) ---------- */
typedef struct {
    int32_t r, c;
    REAL v;
} CAT(Trip, SUF);
#define TRIP CAT(Trip, SUF)

static void FN(csc_free)(CSC* m) {
    free(m->outer);
    free(m->inner);
    free(m->val);
    memset(m, 0, sizeof(*m));
}

static void FN(from_triplets)(CSC* out, int64_t n, const TRIP* t, int64_t nt) {
    FN(csc_free)(out);
    out->n = n;
    int64_t* start = (int64_t*)calloc((size_t)n + 1, sizeof(int64_t));
    for (int64_t i = 0; i < nt; i++) start[t[i].c + 1]++;
    for (int64_t j = 0; j < n; j++) start[j + 1] += start[j];
    /* stable counting sort by column keeps triplet order inside a column */
    int64_t* fill = (int64_t*)malloc(((size_t)n + 1) * sizeof(int64_t));
    memcpy(fill, start, ((size_t)n + 1) * sizeof(int64_t));
    int32_t* rr = (int32_t*)malloc((size_t)(nt ? nt : 1) * sizeof(int32_t));
    REAL* vv = (REAL*)malloc((size_t)(nt ? nt : 1) * sizeof(REAL));
    for (int64_t i = 0; i < nt; i++) {
        int64_t p = fill[t[i].c]++;
        rr[p] = t[i].r;
        vv[p] = t[i].v;
    }
    /* per column: stable insertion sort by row, then collapse (first occurrence + later ones in order) */
    out->outer = (int32_t*)malloc(((size_t)n + 1) * sizeof(int32_t));
    int64_t w = 0;
    out->outer[0] = 0;
    for (int64_t j = 0; j < n; j++) {
        int64_t s = start[j], e = start[j + 1];
        for (int64_t i = s + 1; i < e; i++) {
            int32_t r = rr[i];
            REAL v = vv[i];
            int64_t q = i;
            while (q > s && rr[q - 1] > r) {
                rr[q] = rr[q - 1];
                vv[q] = vv[q - 1];
                q--;
            }
            rr[q] = r;
            vv[q] = v;
        }
        int64_t colstart = w;
        for (int64_t i = s; i < e; i++) {
            if (w > colstart && rr[w - 1] == rr[i]) {
                vv[w - 1] += vv[i];
            } else {
                rr[w] = rr[i];
                vv[w] = vv[i];
                w++;
            }
        }
        out->outer[j + 1] = (int32_t)w;
    }
    out->nnz = w;
    out->inner = (int32_t*)malloc((size_t)(w ? w : 1) * sizeof(int32_t));
    out->val = (REAL*)malloc((size_t)(w ? w : 1) * sizeof(REAL));
    memcpy(out->inner, rr, (size_t)w * sizeof(int32_t));
    memcpy(out->val, vv, (size_t)w * sizeof(REAL));
    free(rr);
    free(vv);
    free(start);
    free(fill);
}

/* FEMContext::update_boundary_conditions (FEMContext.cpp:224-240) + num_unknowns (:46-48) */
static void FN(update_idx_map)(HANDLE* h) {
    int32_t idx = 0;
    for (int64_t i = 0; i < h->V; i++) {
        h->idx_map[i] = -1;
        if (h->bc == 0) {
            if (!h->on_boundary[i]) h->idx_map[i] = idx++;
        } else {
            h->idx_map[i] = idx++;
        }
    }
    h->n = h->V - (h->bc == 0 ? h->num_boundary : 0);
}

/* FEMContext::assemble_{stiffness,mass,advection}_matrix (FEMContext.cpp:70-202) */
static void FN(assemble)(HANDLE* h) {
    int64_t T = h->T;
    TRIP* tk = (TRIP*)malloc((size_t)(9 * T + 1) * sizeof(TRIP));
    TRIP* tm = (TRIP*)malloc((size_t)(9 * T + 1) * sizeof(TRIP));
    TRIP* ta = (TRIP*)malloc((size_t)(9 * T + 1) * sizeof(TRIP));
    int64_t nk = 0, nm = 0, na = 0;
    REAL vel[3] = {(REAL)h->velocity[0], (REAL)h->velocity[1], (REAL)h->velocity[2]};
    for (int64_t k = 0; k < T; k++) {
        ELEM e;
        FN(element)(h, k, &e);
        /* mass: (jdet / 24.0) * {2,1,1;1,2,1;1,1,2}  (:127-132) */
        REAL ms = e.jdet / (REAL)24.0;
        /* advection: t = normalized(v - (v.n)/(|n|*|n|) * n)  (:179-180); Eigen normalized() returns the
         * input unchanged when its squared norm is not > 0 */
        REAL vn = FN(edot3)(vel, e.normal);
        REAL nn = FN(sqrt_)(FN(sum3)(e.normal[0] * e.normal[0], e.normal[1] * e.normal[1], e.normal[2] * e.normal[2]));
        REAL coef = vn / (nn * nn);
        REAL tv[3];
        for (int d = 0; d < 3; d++) tv[d] = vel[d] - coef * e.normal[d];
        REAL tz = FN(sum3)(tv[0] * tv[0], tv[1] * tv[1], tv[2] * tv[2]);
        if (tz > (REAL)0) {
            REAL s = FN(sqrt_)(tz);
            for (int d = 0; d < 3; d++) tv[d] = tv[d] / s;
        }
        REAL as = e.jdet / (REAL)6.0;
        for (int i = 0; i < 3; i++) {
            for (int j = 0; j < 3; j++) {
                uint32_t ni = h->tri[3 * k + i], nj = h->tri[3 * k + j];
                int interior = !h->on_boundary[ni] && !h->on_boundary[nj];
                if (h->bc == 1 || interior) { /* :99, :136 */
                    REAL kv = e.area * FN(edot3)(e.G[i], e.G[j]);
                    tk[nk].r = h->idx_map[ni];
                    tk[nk].c = h->idx_map[nj];
                    tk[nk++].v = kv;
                    tm[nm].r = h->idx_map[ni];
                    tm[nm].c = h->idx_map[nj];
                    tm[nm++].v = ms * (REAL)(i == j ? 2 : 1);
                }
                if (interior) { /* :188 ignores the Neumann flag */
                    ta[na].r = h->idx_map[ni];
                    ta[na].c = h->idx_map[nj];
                    ta[na++].v = as * FN(edot3)(tv, e.G[j]);
                }
            }
        }
    }
    FN(from_triplets)(&h->K, h->n, tk, nk);
    FN(from_triplets)(&h->M, h->n, tm, nm);
    FN(from_triplets)(&h->Adv, h->n, ta, na);
    free(tk);
    free(tm);
    free(ta);
    /* Adv on M's pattern (Adv's pattern is a subset; both sorted) */
    free(h->adv_on_m);
    h->adv_on_m = (REAL*)calloc((size_t)(h->M.nnz ? h->M.nnz : 1), sizeof(REAL));
    for (int64_t j = 0; j < h->n; j++) {
        int64_t p = h->M.outer[j];
        for (int64_t q = h->Adv.outer[j]; q < h->Adv.outer[j + 1]; q++) {
            while (h->M.inner[p] != h->Adv.inner[q]) p++;
            h->adv_on_m[p] = h->Adv.val[q];
        }
    }
}

/* ---------- solver ---------- */
/* y = "A^T x" = row-major walk of the column-major arrays with two accumulators
 * (SparseDenseProduct.h:66-88).  This is what Eigen's CG with Lower|Upper on a col-major matrix
 * multiplies by (ConjugateGradient.h:192-208); equal to A x for the symmetric operators. */
static void FN(spmv_t)(const CSC* P, const REAL* val, const REAL* x, REAL* y) {
    for (int64_t i = 0; i < P->n; i++) {
        REAL ta = 0, tb = 0;
        int64_t k = P->outer[i], e = P->outer[i + 1];
        while (k < e) {
            ta += val[k] * x[P->inner[k]];
            k++;
            if (k < e) {
                tb += val[k] * x[P->inner[k]];
                k++;
            }
        }
        y[i] = ta + tb;
    }
}
/* y = A x exactly as Eigen's col-major sparse * dense: for each column j, y[row] += val * x[j]
 * (SparseDenseProduct.h:96-118); per row this accumulates in ascending column order from 0. */
static void FN(spmv_n)(const CSC* P, const REAL* val, const REAL* x, REAL* y) {
    for (int64_t i = 0; i < P->n; i++) y[i] = 0;
    for (int64_t j = 0; j < P->n; j++) {
        REAL xj = (REAL)1 * x[j];
        for (int64_t k = P->outer[j]; k < P->outer[j + 1]; k++) y[P->inner[k]] += val[k] * xj;
    }
}
/* Dynamic-size dot(): inner_product_impl<Evaluator,true> (lib/eigen/Eigen/src/Core/InnerProduct.h:124-175)
 * as compiled for baseline x86-64 (SSE2 packets of 16 bytes, no FMA in the packet path): four packet
 * accumulators, combined p2+=p3, p1+=p2, p0+=p1, horizontal add (arch/SSE/Reductions.h:205-217:
 * (l0+l2)+(l1+l3); two lanes for double), then the scalar tail with a fused multiply-add. */
#define PK ((int)(16 / sizeof(REAL)))
static REAL FN(predux)(const REAL* l) { return PK == 4 ? (l[0] + l[2]) + (l[1] + l[3]) : l[0] + l[1]; }
static REAL FN(dot)(const REAL* a, const REAL* b, int64_t n) {
    if (n == 0) return 0;
    if (n < PK) {
        REAL r = a[0] * b[0];
        for (int64_t k = 1; k < n; k++) r = FN(fma_)(a[k], b[k], r);
        return r;
    }
    int64_t packetEnd = n - n % PK, quadEnd = n - n % (4 * PK), numPackets = n / PK;
    int64_t numRem = (packetEnd - quadEnd) / PK;
    REAL acc[4][4];
    for (int q = 0; q < 4 && q < numPackets; q++)
        for (int l = 0; l < PK; l++) acc[q][l] = a[q * PK + l] * b[q * PK + l];
    if (numPackets >= 4) {
        for (int64_t k = 4 * PK; k < quadEnd; k += 4 * PK)
            for (int q = 0; q < 4; q++)
                for (int l = 0; l < PK; l++) acc[q][l] = a[k + q * PK + l] * b[k + q * PK + l] + acc[q][l];
        for (int q = 0; q < numRem; q++)
            for (int l = 0; l < PK; l++) acc[q][l] = a[quadEnd + q * PK + l] * b[quadEnd + q * PK + l] + acc[q][l];
        for (int l = 0; l < PK; l++) acc[2][l] = acc[2][l] + acc[3][l];
    }
    if (numPackets >= 3)
        for (int l = 0; l < PK; l++) acc[1][l] = acc[1][l] + acc[2][l];
    if (numPackets >= 2)
        for (int l = 0; l < PK; l++) acc[0][l] = acc[0][l] + acc[1][l];
    REAL r = FN(predux)(acc[0]);
    for (int64_t k = packetEnd; k < n; k++) r = FN(fma_)(a[k], b[k], r);
    return r;
}
/* squaredNorm(): abs2 redux, redux_impl<.., LinearVectorizedTraversal, NoUnrolling>
 * (lib/eigen/Eigen/src/Core/Redux.h:275-324) with two packet accumulators; vectors are 16-byte aligned. */
static REAL FN(sqnorm)(const REAL* a, int64_t n) {
    if (n == 0) return 0;
    int64_t alignedSize2 = (n / (2 * PK)) * (2 * PK), alignedSize = (n / PK) * PK;
    REAL res;
    if (alignedSize) {
        REAL r0[4], r1[4];
        for (int l = 0; l < PK; l++) r0[l] = a[l] * a[l];
        if (alignedSize > PK) {
            for (int l = 0; l < PK; l++) r1[l] = a[PK + l] * a[PK + l];
            for (int64_t i = 2 * PK; i < alignedSize2; i += 2 * PK)
                for (int l = 0; l < PK; l++) {
                    r0[l] = r0[l] + a[i + l] * a[i + l];
                    r1[l] = r1[l] + a[i + PK + l] * a[i + PK + l];
                }
            for (int l = 0; l < PK; l++) r0[l] = r0[l] + r1[l];
            if (alignedSize > alignedSize2)
                for (int l = 0; l < PK; l++) r0[l] = r0[l] + a[alignedSize2 + l] * a[alignedSize2 + l];
        }
        res = FN(predux)(r0);
        for (int64_t i = alignedSize; i < n; i++) res = res + a[i] * a[i];
    } else {
        res = a[0] * a[0];
        for (int64_t i = 1; i < n; i++) res = res + a[i] * a[i];
    }
    return res;
}

static REAL FN(eps)(const HANDLE* h) {
    if (h->tol >= 0) return (REAL)h->tol;
    return sizeof(REAL) == 4 ? (REAL)FLT_EPSILON : (REAL)DBL_EPSILON;
}
static REAL FN(tiny)(void) { return sizeof(REAL) == 4 ? (REAL)FLT_MIN : (REAL)DBL_MIN; }

/* DiagonalPreconditioner::factorize (BasicPreconditioners.h:63-75) */
static void FN(inv_diag)(const CSC* P, const REAL* val, REAL* dinv) {
    for (int64_t j = 0; j < P->n; j++) {
        int64_t k = P->outer[j], e = P->outer[j + 1];
        while (k < e && P->inner[k] != j) k++;
        if (k < e && val[k] != (REAL)0)
            dinv[j] = (REAL)1 / val[k];
        else
            dinv[j] = (REAL)1;
    }
}

/* internal::conjugate_gradient (ConjugateGradient.h:30-85), x0 = 0 (IterativeSolverBase.h:357-360) */
static void FN(pcg)(HANDLE* h, const REAL* aval, const REAL* b, REAL* x, int slot) {
    const CSC* P = &h->M;
    int64_t n = P->n;
    REAL tol = FN(eps)(h);
    int64_t maxit = h->max_iter >= 0 ? h->max_iter : 2 * n;
    REAL* dinv = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    REAL* r = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    REAL* p = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    REAL* z = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    REAL* tmp = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    FN(inv_diag)(P, aval, dinv);
    for (int64_t i = 0; i < n; i++) {
        x[i] = 0;
        r[i] = b[i];
    }
    REAL rhs2 = FN(sqnorm)(b, n);
    int64_t it = 0;
    REAL rn2 = 0;
    if (rhs2 == (REAL)0) {
        h->last_iters[slot] = 0;
        h->last_err[slot] = 0;
        goto done;
    }
    {
        REAL thr = tol * tol * rhs2;
        if (thr < FN(tiny)()) thr = FN(tiny)();
        rn2 = FN(sqnorm)(r, n);
        if (rn2 < thr) goto fin;
        for (int64_t i = 0; i < n; i++) p[i] = dinv[i] * r[i];
        REAL absNew = FN(dot)(r, p, n);
        while (it < maxit) {
            FN(spmv_t)(P, aval, p, tmp);
            REAL alpha = absNew / FN(dot)(p, tmp, n);
            for (int64_t i = 0; i < n; i++) x[i] += alpha * p[i];
            for (int64_t i = 0; i < n; i++) r[i] -= alpha * tmp[i];
            rn2 = FN(sqnorm)(r, n);
            if (rn2 < thr) break;
            for (int64_t i = 0; i < n; i++) z[i] = dinv[i] * r[i];
            REAL absOld = absNew;
            absNew = FN(dot)(r, z, n);
            REAL beta = absNew / absOld;
            for (int64_t i = 0; i < n; i++) p[i] = z[i] + beta * p[i];
            it++;
        }
    }
fin:
    h->last_iters[slot] = it;
    h->last_err[slot] = sqrt((double)(rn2 / rhs2));
done:
    free(dinv);
    free(r);
    free(p);
    free(z);
    free(tmp);
}

/* Jacobi right-preconditioned BiCGSTAB following internal::bicgstab (BiCGSTAB.h:31-119), x0 = 0,
 * multiplying by A (column walk, spmv_n).  The reference never instantiates BiCGSTAB (fact: every
 * equation uses ConjugateGradient, CPUSolver.cpp:65); this is the "correct non-symmetric mode" the
 * north star adds.  Two stated deviations from the vendored file: the stopping test is relative,
 * |r| > tol*|b| (the vendored dev snapshot compares |r| with tol directly), and the small-angle
 * breakdown restarts with r0 = r instead of r0.setRandom() so that results are reproducible. */
static REAL FN(norm)(const REAL* a, int64_t n) { return FN(sqrt_)(FN(sqnorm)(a, n)); }
static void FN(bicgstab)(HANDLE* h, const REAL* aval, const REAL* b, REAL* x, int slot) {
    const CSC* P = &h->M;
    int64_t n = P->n;
    REAL tol = FN(eps)(h);
    int64_t maxit = h->max_iter >= 0 ? h->max_iter : 2 * n;
    size_t sz = (size_t)(n + 1) * sizeof(REAL);
    REAL *dinv = malloc(sz), *r = malloc(sz), *r0 = malloc(sz), *v = malloc(sz), *p = malloc(sz), *y = malloc(sz),
         *z = malloc(sz), *s = malloc(sz), *t = malloc(sz);
    FN(inv_diag)(P, aval, dinv);
    for (int64_t i = 0; i < n; i++) {
        x[i] = 0;
        r[i] = b[i];
        r0[i] = b[i];
        v[i] = 0;
        p[i] = 0;
    }
    REAL r0_norm = FN(norm)(r0, n);
    REAL r_norm = r0_norm;
    REAL rhs_norm = FN(norm)(b, n);
    int64_t i = 0, restarts = 0;
    if (rhs_norm == (REAL)0) {
        h->last_iters[slot] = 0;
        h->last_err[slot] = 0;
        goto done;
    }
    {
        REAL rho = 1, alpha = 0, w = 1;
        REAL eps = (sizeof(REAL) == 4 ? (REAL)FLT_EPSILON : (REAL)DBL_EPSILON);
        while (r_norm > tol * rhs_norm && i < maxit) {
            REAL rho_old = rho;
            rho = FN(dot)(r0, r, n);
            REAL mx = r0_norm > r_norm ? r0_norm : r_norm, mn = r0_norm < r_norm ? r0_norm : r_norm;
            if ((REAL)fabs((double)rho) / mx < eps * mn) {
                FN(spmv_n)(P, aval, x, t);
                for (int64_t q = 0; q < n; q++) {
                    r[q] = b[q] - t[q];
                    r0[q] = r[q];
                }
                rho = FN(sqnorm)(r, n);
                r0_norm = FN(norm)(r, n);
                alpha = 0;
                w = 1;
                if (restarts++ == 0) i = 0;
            }
            REAL beta = (rho / rho_old) * (alpha / w);
            for (int64_t q = 0; q < n; q++) p[q] = r[q] + beta * (p[q] - w * v[q]);
            for (int64_t q = 0; q < n; q++) y[q] = dinv[q] * p[q];
            FN(spmv_n)(P, aval, y, v);
            REAL theta = FN(dot)(r0, v, n);
            REAL v_norm = FN(norm)(v, n);
            mx = r0_norm > v_norm ? r0_norm : v_norm;
            mn = r0_norm < v_norm ? r0_norm : v_norm;
            if ((REAL)fabs((double)theta) / mx < eps * mn) {
                FN(spmv_n)(P, aval, x, t);
                for (int64_t q = 0; q < n; q++) {
                    r[q] = b[q] - t[q];
                    r0[q] = r[q];
                }
                r0_norm = FN(norm)(r0, n);
                r_norm = r0_norm;
                rho = 1;
                alpha = 0;
                w = 1;
                if (restarts++ == 0) i = 0;
                if (restarts > 8) break;
                continue;
            }
            alpha = rho / theta;
            for (int64_t q = 0; q < n; q++) s[q] = r[q] - alpha * v[q];
            for (int64_t q = 0; q < n; q++) z[q] = dinv[q] * s[q];
            FN(spmv_n)(P, aval, z, t);
            REAL tmp = FN(sqnorm)(t, n);
            if (tmp > (REAL)0)
                w = FN(dot)(t, s, n) / tmp;
            else
                w = 0;
            for (int64_t q = 0; q < n; q++) x[q] += alpha * y[q] + w * z[q];
            for (int64_t q = 0; q < n; q++) r[q] = s[q] - w * t[q];
            r_norm = FN(norm)(r, n);
            ++i;
        }
    }
    h->last_iters[slot] = i;
    h->last_err[slot] = (double)(r_norm / rhs_norm);
done:
    free(dinv);
    free(r);
    free(r0);
    free(v);
    free(p);
    free(y);
    free(z);
    free(s);
    free(t);
}

/* CPUSolver::get_surface_value_vector / map_vector_to_surface (CPUSolver.cpp:13-36) */
static void FN(gather)(const HANDLE* h, REAL* x) {
    for (int64_t i = 0; i < h->V; i++)
        if (h->idx_map[i] != -1) x[h->idx_map[i]] = (REAL)h->values[i];
}
static void FN(scatter)(HANDLE* h, const REAL* x) {
    for (int64_t i = 0; i < h->V; i++) h->values[i] = h->idx_map[i] != -1 ? (float)x[h->idx_map[i]] : 0.0f;
}

/* CPUSolver::advance_time (CPUSolver.cpp:64-147).  Scalars are the reference's floats promoted to REAL;
 * products are formed in the reference's association order. */
static void FN(advance)(HANDLE* h) {
    int64_t n = h->n, nnz = h->M.nnz;
    const CSC* P = &h->M;
    const float* q = h->par[h->equation];
    REAL dt = (REAL)q[0];
    REAL* A = (REAL*)malloc((size_t)(nnz + 1) * sizeof(REAL));
    REAL* Md = (REAL*)malloc((size_t)(nnz + 1) * sizeof(REAL));
    REAL* b = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    REAL* b2 = (REAL*)malloc((size_t)(n + 1) * sizeof(REAL));
    for (int64_t k = 0; k < nnz; k++) Md[k] = h->M.val[k] / dt;
    h->last_iters[0] = h->last_iters[1] = 0;
    h->last_err[0] = h->last_err[1] = 0;
    switch (h->equation) {
        case 0: { /* Heat :72-84 */
            REAL kap = (REAL)q[1];
            FN(gather)(h, h->u);
            for (int64_t k = 0; k < nnz; k++) A[k] = Md[k] + kap * h->K.val[k];
            FN(spmv_n)(P, Md, h->u, b);
            FN(pcg)(h, A, b, h->u, 0);
            FN(scatter)(h, h->u);
        } break;
        case 2: { /* Advection-diffusion :90-102 */
            REAL c = (REAL)q[1];
            FN(gather)(h, h->u);
            for (int64_t k = 0; k < nnz; k++) A[k] = (Md[k] + c * h->K.val[k]) - h->adv_on_m[k];
            FN(spmv_n)(P, Md, h->u, b);
            if (h->solver_mode == 1)
                FN(bicgstab)(h, A, b, h->u, 0);
            else
                FN(pcg)(h, A, b, h->u, 0);
            FN(scatter)(h, h->u);
        } break;
        case 1: { /* Wave :108-121 */
            REAL c = (REAL)q[1];
            REAL cc = c * c;
            FN(gather)(h, h->u);
            for (int64_t k = 0; k < nnz; k++) A[k] = Md[k] + (cc * h->K.val[k]) * dt;
            REAL* ck = (REAL*)malloc((size_t)(nnz + 1) * sizeof(REAL));
            for (int64_t k = 0; k < nnz; k++) ck[k] = cc * h->K.val[k];
            /* "P1 - P2" of two products is evaluated as dst = P1; dst -= P2, the second through
             * scaleAndAddTo(alpha = -1): rhs_j = -1 * u_j and dst[row] += (cc K_row,j) * rhs_j, i.e. the
             * K.u terms keep accumulating onto the M.v sum in ascending column order
             * (ProductEvaluators.h assignment_from_xpr_op_product, SparseDenseProduct.h:96-118). */
            FN(spmv_n)(P, Md, h->v, b);
            for (int64_t j = 0; j < n; j++) {
                REAL xj = (REAL)(-1) * h->u[j];
                for (int64_t k = P->outer[j]; k < P->outer[j + 1]; k++) b[P->inner[k]] += ck[k] * xj;
            }
            free(ck);
            FN(pcg)(h, A, b, h->v, 0);
            for (int64_t i = 0; i < n; i++) h->u[i] = h->u[i] + h->v[i] * dt;
            FN(scatter)(h, h->u);
        } break;
        default: { /* Gray-Scott :127-145 */
            REAL Du = (REAL)q[1], Dv = (REAL)q[2], f = (REAL)q[3], kk = (REAL)q[4];
            REAL fk = f + kk;
            REAL* A2 = (REAL*)malloc((size_t)(nnz + 1) * sizeof(REAL));
            FN(gather)(h, h->v);
            for (int64_t k = 0; k < nnz; k++) A[k] = Md[k] + Du * h->K.val[k];
            for (int64_t k = 0; k < nnz; k++) A2[k] = Md[k] + Dv * h->K.val[k];
            FN(spmv_n)(P, Md, h->u, b);
            FN(spmv_n)(P, Md, h->v, b2);
            for (int64_t i = 0; i < n; i++) {
                REAL uvv = h->u[i] * (h->v[i] * h->v[i]);
                b[i] = (b[i] - uvv) + f * ((REAL)1 - h->u[i]);
                b2[i] = (b2[i] + uvv) - fk * h->v[i];
            }
            FN(pcg)(h, A, b, h->u, 0);
            FN(pcg)(h, A2, b2, h->v, 1);
            FN(scatter)(h, h->v);
            free(A2);
        } break;
    }
    free(A);
    free(Md);
    free(b);
    free(b2);
}

/* ---------- C ABI (ctypes) ---------- */
void* FN(create)(int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* on_boundary, int bc) {
    HANDLE* h = (HANDLE*)calloc(1, sizeof(HANDLE));
    h->V = V;
    h->T = T;
    h->xyz = (float*)malloc((size_t)(3 * V + 1) * sizeof(float));
    memcpy(h->xyz, xyz, (size_t)(3 * V) * sizeof(float));
    h->tri = (uint32_t*)malloc((size_t)(3 * T + 1) * sizeof(uint32_t));
    memcpy(h->tri, tri, (size_t)(3 * T) * sizeof(uint32_t));
    h->on_boundary = (uint8_t*)malloc((size_t)V + 1);
    for (int64_t i = 0; i < V; i++) {
        h->on_boundary[i] = on_boundary[i] != 0;
        h->num_boundary += h->on_boundary[i];
    }
    h->bc = bc;
    h->equation = 1; /* FEMContext.hpp:31 default Wave */
    /* EquationParameters.hpp defaults */
    float d[4][5] = {{0.01f, 0.05f, 0, 0, 0}, {0.05f, 0.05f, 0, 0, 0}, {0.001f, 0.25f, 0, 0, 0}, {0.001f, 0.08f, 0.04f, 0.035f, 0.06f}};
    memcpy(h->par, d, sizeof(d));
    h->velocity[0] = 1.0f;
    h->tol = -1;
    h->max_iter = -1;
    h->idx_map = (int32_t*)malloc((size_t)(V + 1) * sizeof(int32_t));
    h->values = (float*)calloc((size_t)V + 1, sizeof(float));
    FN(update_idx_map)(h);
    FN(assemble)(h);
    h->u = (REAL*)calloc((size_t)h->n + 1, sizeof(REAL));
    h->v = (REAL*)calloc((size_t)h->n + 1, sizeof(REAL));
    return h;
}
void FN(destroy)(void* p) {
    HANDLE* h = (HANDLE*)p;
    FN(csc_free)(&h->M);
    FN(csc_free)(&h->K);
    FN(csc_free)(&h->Adv);
    free(h->adv_on_m);
    free(h->xyz);
    free(h->tri);
    free(h->on_boundary);
    free(h->idx_map);
    free(h->values);
    free(h->u);
    free(h->v);
    free(h);
}
static void FN(reset_state)(HANDLE* h) { /* CPUSolver::clear_values :54-59 */
    free(h->u);
    free(h->v);
    h->u = (REAL*)calloc((size_t)h->n + 1, sizeof(REAL));
    h->v = (REAL*)calloc((size_t)h->n + 1, sizeof(REAL));
}
void FN(set_boundary_condition)(void* p, int bc) {
    HANDLE* h = (HANDLE*)p;
    h->bc = bc;
    FN(update_idx_map)(h);
    FN(assemble)(h);
    FN(reset_state)(h);
}
void FN(set_equation)(void* p, int eq) { ((HANDLE*)p)->equation = eq; }
void FN(set_params)(void* p, int eq, const float* q) { memcpy(((HANDLE*)p)->par[eq], q, 5 * sizeof(float)); }
void FN(set_velocity)(void* p, const float* vel) {
    HANDLE* h = (HANDLE*)p;
    memcpy(h->velocity, vel, 3 * sizeof(float));
    FN(assemble)(h);
}
void FN(set_solver)(void* p, int mode, double tol, int64_t max_iter) {
    HANDLE* h = (HANDLE*)p;
    h->solver_mode = mode;
    h->tol = tol;
    h->max_iter = max_iter;
}
int64_t FN(num_nodes)(void* p) { return ((HANDLE*)p)->V; }
int64_t FN(num_unknowns)(void* p) { return ((HANDLE*)p)->n; }
int FN(max_row_nonzeros)(void* p) { /* FEMContext::compute_max_row_nonzeros :207-219 (on K) */
    HANDLE* h = (HANDLE*)p;
    int m = 0;
    for (int64_t j = 0; j < h->n; j++) {
        int c = h->K.outer[j + 1] - h->K.outer[j];
        if (c > m) m = c;
    }
    return m;
}
void FN(get_idx_map)(void* p, int32_t* out) { memcpy(out, ((HANDLE*)p)->idx_map, (size_t)((HANDLE*)p)->V * sizeof(int32_t)); }
static CSC* FN(pick)(HANDLE* h, int which) { return which == 0 ? &h->M : which == 1 ? &h->K : &h->Adv; }
int64_t FN(nnz)(void* p, int which) { return FN(pick)((HANDLE*)p, which)->nnz; }
void FN(get_csc)(void* p, int which, int32_t* outer, int32_t* inner, REAL* val) {
    CSC* m = FN(pick)((HANDLE*)p, which);
    memcpy(outer, m->outer, (size_t)(m->n + 1) * sizeof(int32_t));
    memcpy(inner, m->inner, (size_t)m->nnz * sizeof(int32_t));
    memcpy(val, m->val, (size_t)m->nnz * sizeof(REAL));
}
void FN(clear)(void* p) {
    HANDLE* h = (HANDLE*)p;
    FN(reset_state)(h);
    memset(h->values, 0, (size_t)h->V * sizeof(float));
}
void FN(set_values)(void* p, const float* v) { memcpy(((HANDLE*)p)->values, v, (size_t)((HANDLE*)p)->V * sizeof(float)); }
void FN(get_values)(void* p, float* v) { memcpy(v, ((HANDLE*)p)->values, (size_t)((HANDLE*)p)->V * sizeof(float)); }
void FN(get_state)(void* p, int which, REAL* out) {
    HANDLE* h = (HANDLE*)p;
    memcpy(out, which == 0 ? h->u : h->v, (size_t)h->n * sizeof(REAL));
}
/* test hook: overwrite hidden state (used to start the GPU path and the oracle from the same point) */
void FN(set_state)(void* p, int which, const REAL* in) {
    HANDLE* h = (HANDLE*)p;
    memcpy(which == 0 ? h->u : h->v, in, (size_t)h->n * sizeof(REAL));
}
void FN(step)(void* p, int nsteps) {
    for (int i = 0; i < nsteps; i++) FN(advance)((HANDLE*)p);
}
int FN(unstable)(void* p) { /* CPUSolver::has_numerical_instability :41-48 (signed compare, no abs) */
    HANDLE* h = (HANDLE*)p;
    for (int64_t i = 0; i < h->n; i++)
        if (h->u[i] > (REAL)1e4f || h->v[i] > (REAL)1e4f) return 1;
    return 0;
}
void FN(last_stats)(void* p, int64_t* iters, double* err) {
    HANDLE* h = (HANDLE*)p;
    iters[0] = h->last_iters[0];
    iters[1] = h->last_iters[1];
    err[0] = h->last_err[0];
    err[1] = h->last_err[1];
}

#undef PK
#undef CSC
#undef ELEM
#undef TRIP
#undef HANDLE
#undef FN
