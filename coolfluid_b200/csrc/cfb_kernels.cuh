// cfb200 — the FVMCC ComputeRHS / ComputeRhsJacob kernels, written as per-item functors
// (`operator()(int idx)`: one cell, node, face or face-side per call) over the device
// structure-of-arrays `Dev`.  cfb_engine.cu launches each functor as a CUDA kernel with one
// thread per item.  Design:
//   * no floating-point atomics anywhere: faces write their flux / Jacobian blocks to face-owned
//     slots, cells GATHER them in the reference's face-loop order, so every sum is formed in exactly
//     the order of the sequential CPU loop (FVMCC_ComputeRHS.cxx:202-319) and is run-to-run deterministic;
//   * off-diagonal Jacobian blocks have exactly one writer (the face between the two cells) and are
//     stored straight into the BSR value array; the diagonal block of a cell is minus the sum of the
//     blocks in its COLUMN (block(n,c) for each face neighbour n) plus its boundary-face blocks, again
//     gathered in face order (FVMCC_ComputeRhsJacob.cxx:347-365: row R receives -1 x what row L receives);
//   * all reals FP64, SoA by component so that one-thread-per-item accesses coalesce.
#pragma once
#define CFB_SA_WIDTH 6     /* faces per cell listed for RoeSA: hexahedra, prisms, pyramids, tetrahedra, quads, triangles */
#include "cfb_math.cuh"

namespace cfb {

struct Dev {
  int dim, nCells, nGhosts, nStates, nNodes, nFaces, nInner;
  // ---- static mesh data (uploaded once) ----
  const double* nodes;        // [d][nNodes]
  const double* centres;      // [d][nStates]   cells then ghosts
  const int* faceL; const int* faceR;
  const int* faceNodePtr; const int* faceNodes;
  const double* normals;      // [d][nFaces]
  const double* areas;        // [nFaces]
  const double* faceCentres;  // [d][nFaces]    face quadrature point = node average (FVMCC_PolyRec.hh:114-123)
  const double* volumes;
  const int* cellFacePtr; const int* cellFaces;    // reference local face order (limiter loop)
  const int* cellFaceSorted;  // same CSR; faces ascending in loop order; ~f when the cell is the face's R state
  const int* stencilPtr; const int* stencil;
  const int* lsqPtr; const int* lsqNbr;            // per-cell neighbour list in the reference's edge order
  const int* nodeNbrPtr; const int* nodeNbr; const double* nodeW;
  const double* cvol; const uint8_t* cvSign; const int* faceBC;
  const double* bcFaceVal;    // [q][nGhosts] position-dependent Dirichlet data (cfb200_set_bc_face_values), or null
  const uint8_t* bcFaceFlag;  // [nGhosts] 1 where bcFaceVal replaces BCond::value
  const double* bcWallRatio;  // [nGhosts] (inner-ghost)/(inner-wall) distance ratio of FVMCC_BaseBC::computeGhostPosition, or null
  const int* bndStPtr; const int* bndSt;           // Corrected2D/3D: gradient stencil of each boundary face (CSR over ghosts)
  const int* ggPtr; const int* ggNode;   // CorrectedGG: per cell, the nodes of its faces in computeGradientsGG order (CSR over cells)
  const int* ggAv;            // CorrectedGG: [4][nFaces] entries of the "average state": node id >= 0, ~state id < 0, INT32_MIN none
  const int* saFace;          // "RoeSA": [2*CFB_SA_WIDTH][nFaces] faces of the L / R cell entering etaSA (-1 = none), or null
  const uint8_t* updatable;
  const uint8_t* cellActive;  // cell touches at least one face with an updatable side
  const BCond* bcs;
  // storage permutation (cfg.renumber): reference (caller) ID of internal cell i; null = the identity.  rhs, updateCoeff,
  // the norm partials and the BSR rows are indexed by REFERENCE id (they are the consumer's), everything else by storage id
  const int* cellRef;
  double* lsqL;               // [6][nCells]  l11 l12 l13 l22 l23 l33
  double* lsqW;               // [len(lsqNbr)] inverse-distance weight of each gather entry (static geometry: formed once at setup)
  // ---- block Jacobian ----
  const int* bsrRowPtr; const int* bsrColInd;
  const int* slotLR;          // [nInner] block (L,R)
  const int* slotRL;          // [nInner] block (R,L)
  const int* diagSlot;        // [nCells]
  const int* diagSrc;         // CSR like cellFaceSorted: source block of each face's diagonal contribution (see DiagGather)
  double* bsrVal;             // [nnzb][q][q]
  double* bndBlocks;          // [nGhosts][q][q]  boundary-face contribution to the diagonal of the inner cell
  // ---- fields ----
  double* U;                  // [q][nStates]
  double* pastU;              // [q][nCells]
  double* uX; double* uY; double* uZ; double* lim;   // [q][nCells]
  double* nstates;            // [q][nNodes]
  double* faceFlux;           // [q][nFaces]  resFactor is NOT applied here
  double* ucL; double* ucR; double* ucD; double* faceMu;   // [nFaces]
  double* rhs;                // AoS [nCells][q] (consumer layout: PETSc/Paralution block vectors)
  double* updateCoeff;        // [nCells]
  double* aos;                // staging [max(nStates,nNodes)*q]
  double* partial;            // reduction scratch
  int* status;
  // "DistanceBasedGMove" (DistanceBasedExtrapolatorGMove.cxx:167-173): nodes whose extrapolated values nodePrescIdx[i] are
  // overwritten by nodePrescVal[i] (cfb200_set_nodal_prescribed); null = plain "DistanceBased".  Kept at the END of the
  // struct: the kernel-parameter offsets of everything above are those of the binaries profiled in round 2 (the face kernel's
  // register allocation moved from 252 to 238 registers, 27.4 -> 28.3 ms, when these sat in the middle)
  const uint8_t* nodePresc; const int* nodePrescIdx; const double* nodePrescVal; int nNodePresc;
};

CFB_HD int refId(const Dev& d, int cell) { return d.cellRef ? d.cellRef[cell] : cell; }

// ------------------------------------------------------------------------------------------------
// layout conversion between the reference's AoS sockets and the device SoA; ref (optional): storage item i is the
// caller's item ref[i] (cells / faces renumbered inside the engine)
struct AosToSoa {   // idx over n*q
  const double* src; double* dst; int n, q, ld; const int* ref;
  CFB_HD void operator()(int idx) const {
    const int i = idx/q, v = idx - i*q;
    dst[(size_t)v*ld + i] = src[ref ? (size_t)ref[i]*q + v : (size_t)idx];
  }
};
struct SoaToAos {
  const double* src; double* dst; int n, q, ld; const int* ref;
  CFB_HD void operator()(int idx) const {
    const int i = idx/q, v = idx - i*q;
    dst[ref ? (size_t)ref[i]*q + v : (size_t)idx] = src[(size_t)v*ld + i];
  }
};
struct FaceCentres {   // idx over faces
  Dev d;
  CFB_HD void operator()(int f) const {
    const int b = d.faceNodePtr[f], e = d.faceNodePtr[f+1];
    const double ov = 1./(e - b);
    for (int k = 0; k < d.dim; ++k) {
      double x = 0.0;
      for (int j = b; j < e; ++j) x += d.nodes[(size_t)k*d.nNodes + d.faceNodes[j]]*ov;
      const_cast<double*>(d.faceCentres)[(size_t)k*d.nFaces + f] = x;
    }
  }
};

// ------------------------------------------------------------------------------------------------
// a16: boundary conditions, FVMCC_BaseBC.ci:144-187 executeOnTrs -> setGhostState(face)
// g = boundary face (ghost) index: position-dependent Dirichlet data, when set, replaces the constant of the record
// (SuperInlet.cxx:156-209 evaluates its Def function at the boundary point of every face)
template <int D, class PH = Phys<D> >
CFB_HD void ghostState(const Dev& d, int g, const Cfg& c, const BCond& b, const double* nr /*area-weighted normal*/,
                       const double* inner, double* ghost) {
  constexpr int Q = PH::Q;
  typedef PH P;
  switch (b.type) {
    case CFB200_BC_SUPER_OUTLET:
      for (int i = 0; i < Q; ++i) ghost[i] = inner[i];
      break;
    case CFB200_BC_DIRICHLET:
      if (d.bcFaceFlag && d.bcFaceFlag[g]) {
        for (int i = 0; i < Q; ++i)
          ghost[i] = b.mask[i] ? (2.*d.bcFaceVal[(size_t)i*d.nGhosts + g] - inner[i]) : inner[i];
      } else {
        for (int i = 0; i < Q; ++i) ghost[i] = b.mask[i] ? (2.*b.value[i] - inner[i]) : inner[i];
      }
      break;
    case CFB200_BC_MIRROR_VELOCITY: {     // MirrorVelocity.cxx:71-111
      double vn = 0., area2 = 0.;
      for (int i = 0; i < D; ++i) { const double nxComp = nr[i]; vn += nxComp*inner[1+i]; area2 += nxComp*nxComp; }
      ghost[0] = inner[0];
      for (int i = 0; i < D; ++i) ghost[1+i] = inner[1+i] - 2.0*vn*nr[i]/area2;
      ghost[Q-1] = inner[Q-1];
      break;
    }
    case CFB200_BC_FARFIELD_EULER:        // FarFieldEuler2D.cxx:127-300, FarFieldEuler3D.cxx:62-196
      P::farFieldGhost(c, nr, b.value, inner, ghost);
      break;
    case CFB200_BC_NOSLIP_ISOTHERMAL: {   // NoSlipWallIsothermalNSvt.ci:178-243: linear extrapolation through the wall values
      constexpr int TID = Q - 1;
      const double ratio = d.bcWallRatio[g];                     // m_drXiXg/m_drXiXw (static geometry: formed at set_mesh)
      const double ghostT = 2.*b.value[TID] - inner[TID];        // repositionNode: ghostValue = 2 bValue - innerValue
      if (!(ghostT >= b.value[0])) *d.status = CFB200_ERR_GHOST_REPOSITION;      // the reference would move the ghost node
      for (int i = 0; i < Q; ++i) {
        if (i >= 1 && i <= D) ghost[i] = inner[i] - (inner[i] - b.value[i])*ratio;
        else if (i != TID) ghost[i] = inner[i];
        else ghost[i] = ghostT;
      }
      break;
    }
    case CFB200_BC_FARFIELD_EULER_CHAR:   // FarFieldEulerChar2D.cxx:36-254, FarFieldEulerChar3D.cxx:40-200
      if constexpr (Q != 9) P::farFieldCharGhost(c, nr, b.value, inner, ghost);
      else { for (int i = 0; i < Q; ++i) ghost[i] = inner[i]; }      // rejected at set_mesh for MHD
      break;
    case CFB200_BC_SUPER_OUTLET_MHD:      // SuperOutletMHD3DProjection.cxx:76-239: reads the inner and ghost state coordinates
      if constexpr (Q == 9) {
        const int L = d.faceL[d.nInner + g];
        double xi[3], xg[3];
        for (int k = 0; k < 3; ++k) {
          xi[k] = d.centres[(size_t)k*d.nStates + L];
          xg[k] = d.centres[(size_t)k*d.nStates + d.nCells + g];
        }
        P::superOutletGhost(c, b.mask[0] != 0, b.value[0], xi, xg, inner, ghost);
      } else {
        for (int i = 0; i < Q; ++i) ghost[i] = inner[i];       // rejected at set_mesh for the Euler models
      }
      break;
    default:                              // CFB200_BC_MIRROR_EULER_CONS: MirrorEuler2D.cxx:56-94 / MirrorMHD3DProjection.cxx:56-96
      P::mirrorGhost(c, nr, inner, ghost);
      break;
  }
}

template <int D, class PH = Phys<D> >
struct GhostFill {   // idx over boundary faces
  Dev d; Cfg c;
  CFB_HD void operator()(int g) const {
    constexpr int Q = PH::Q;
    const int f = d.nInner + g;
    const int L = d.faceL[f];
    double inner[Q], ghost[Q], nr[D];
    for (int v = 0; v < Q; ++v) inner[v] = d.U[(size_t)v*d.nStates + L];
    for (int k = 0; k < D; ++k) nr[k] = d.normals[(size_t)k*d.nFaces + f];
    ghostState<D, PH>(d, g, c, d.bcs[d.faceBC[f]], nr, inner, ghost);
    for (int v = 0; v < Q; ++v) d.U[(size_t)v*d.nStates + d.nCells + g] = ghost[v];
  }
};

// ------------------------------------------------------------------------------------------------
// a4/a5: least-squares gradients.  The reference accumulates edge by edge, each edge visited once from
// its lower-ID cell (LeastSquareP1PolyRec2D.cxx:100-134, 3D.cxx:104-160).  `lsqNbr` lists, per cell, the
// other end of every edge that touches it in exactly that global edge order, so the per-cell gather below
// forms the same sums term by term ((-dx)*(-du) == dx*du exactly, so the edge orientation is irrelevant).
template <int D, class PH = Phys<D> >
CFB_HD double edgeWeight(const Dev& d, int c, int nb, double* dx) {
  double dist = 0.;
  for (int k = 0; k < D; ++k) {
    const double xc = d.centres[(size_t)k*d.nStates + c], xn = d.centres[(size_t)k*d.nStates + nb];
    const double diff = xc - xn;
    dist += diff*diff;
    dx[k] = xn - xc;
  }
  const double w = ddiv(1.0, dsqrt(dist));
  for (int k = 0; k < D; ++k) dx[k] = w*dx[k];
  return w;
}
// the same with the weight read back from the setup pass (LeastSquareP1PolyRec2D::updateWeights keeps them in its
// `weights` array too, LeastSquareP1PolyRec2D.cxx:266-400): no square root / division per neighbour and evaluation
template <int D, class PH = Phys<D> >
CFB_HD double edgeWeightStored(const Dev& d, int c, int nb, int j, double* dx) {
  const double w = d.lsqW[j];
  for (int k = 0; k < D; ++k) {
    const double xc = d.centres[(size_t)k*d.nStates + c], xn = d.centres[(size_t)k*d.nStates + nb];
    dx[k] = w*(xn - xc);
  }
  return w;
}

template <int D, class PH = Phys<D> >
struct LsqSetup {   // LeastSquareP1PolyRec{2,3}D::updateWeights; idx over cells
  Dev d;
  CFB_HD void operator()(int c) const {
    double l[6] = {0., 0., 0., 0., 0., 0.};
    for (int j = d.lsqPtr[c]; j < d.lsqPtr[c+1]; ++j) {
      double dx[3] = {0., 0., 0.};
      d.lsqW[j] = edgeWeight<D>(d, c, d.lsqNbr[j], dx);
      if (D == 2) { l[0] += dx[0]*dx[0]; l[1] += dx[0]*dx[1]; l[3] += dx[1]*dx[1]; }
      else {
        l[0] += dx[0]*dx[0]; l[1] += dx[0]*dx[1]; l[2] += dx[0]*dx[2];
        l[3] += dx[1]*dx[1]; l[4] += dx[1]*dx[2]; l[5] += dx[2]*dx[2];
      }
    }
    for (int k = 0; k < 6; ++k) d.lsqL[(size_t)k*d.nCells + c] = l[k];
  }
};

template <int D, class PH = Phys<D> >
struct Gradients {   // idx over the cells [c0, c0 + n), or over the cell list `cells` when it is given
  Dev d; int c0; const int* cells;
  CFB_HD void operator()(int idx) const {
    constexpr int Q = PH::Q;
    const int c = cells ? cells[idx] : c0 + idx;
    double lf[3][Q];
    double u0[Q];
    for (int v = 0; v < Q; ++v) { lf[0][v] = lf[1][v] = lf[2][v] = 0.; u0[v] = d.U[(size_t)v*d.nStates + c]; }
    // (batching the gathers of several neighbours was measured: 142 registers, 2.07 instead of 1.97 ms at 10.6 M cells)
    for (int j = d.lsqPtr[c]; j < d.lsqPtr[c+1]; ++j) {
      const int nb = d.lsqNbr[j];
      double dx[3] = {0., 0., 0.};
      const double w = edgeWeightStored<D>(d, c, nb, j, dx);
      for (int v = 0; v < Q; ++v) {
        const double du = w*(d.U[(size_t)v*d.nStates + nb] - u0[v]);
        lf[0][v] += dx[0]*du; lf[1][v] += dx[1]*du;
        if (D == 3) lf[2][v] += dx[2]*du;
      }
    }
    if (D == 2) {
      const double l11 = d.lsqL[c], l12 = d.lsqL[(size_t)d.nCells + c], l22 = d.lsqL[(size_t)3*d.nCells + c];
      const double invDet = ddiv(1., l11*l22 - l12*l12);
      for (int v = 0; v < Q; ++v) {
        d.uX[(size_t)v*d.nCells + c] = (l22*lf[0][v] - l12*lf[1][v])*invDet;
        d.uY[(size_t)v*d.nCells + c] = (l11*lf[1][v] - l12*lf[0][v])*invDet;
      }
    } else {
      const size_t n = d.nCells;
      const double l11 = d.lsqL[c], l12 = d.lsqL[n + c], l13 = d.lsqL[2*n + c];
      const double l22 = d.lsqL[3*n + c], l23 = d.lsqL[4*n + c], l33 = d.lsqL[5*n + c];
      const double det = l11*l22*l33 - l11*l23*l23 - l12*l12*l33 + l12*l13*l23 + l13*l12*l23 - l13*l13*l22;
      const double linv11 = l22*l33 - l23*l23;
      const double linv22 = l11*l33 - l13*l13;
      const double linv33 = l11*l22 - l12*l12;
      const double linv12 = -(l12*l33 - l13*l23);
      const double linv13 = l12*l23 - l13*l22;
      const double linv23 = -(l11*l23 - l13*l12);
      const bool ok = !(fabs(det) <= 2.2250738585072014e-308);   // !MathChecks::isZero(det)
      const Recip rdet = recipOf(ok ? det : 1.0);
      for (int v = 0; v < Q; ++v) {
        double gx = 0.0, gy = 0.0, gz = 0.0;
        if (ok) {
          gx = divBy(linv11*lf[0][v] + linv12*lf[1][v] + linv13*lf[2][v], rdet);
          gy = divBy(linv12*lf[0][v] + linv22*lf[1][v] + linv23*lf[2][v], rdet);
          gz = divBy(linv13*lf[0][v] + linv23*lf[1][v] + linv33*lf[2][v], rdet);
        }
        d.uX[(size_t)v*n + c] = gx; d.uY[(size_t)v*n + c] = gy; d.uZ[(size_t)v*n + c] = gz;
      }
    }
  }
};

// ------------------------------------------------------------------------------------------------
// a6: nodal states, DistanceBasedExtrapolator.ci:105-149; idx over nodes
template <int D, class PH = Phys<D> >
struct NodalExtrapolate {
  Dev d;
  CFB_HD void operator()(int n) const {
    constexpr int Q = PH::Q;
    double ns[Q];
    for (int v = 0; v < Q; ++v) ns[v] = 0.;
    for (int k = d.nodeNbrPtr[n]; k < d.nodeNbrPtr[n+1]; ++k) {
      const double w = d.nodeW[k];
      const int s = d.nodeNbr[k];
      for (int v = 0; v < Q; ++v) ns[v] += d.U[(size_t)v*d.nStates + s]*w;
    }
    if (d.nodePresc && d.nodePresc[n]) {            // applyBC() after applyInner(), DistanceBasedExtrapolatorGMove.cxx:140-173
      for (int i = 0; i < d.nNodePresc; ++i) {
        const int idx = d.nodePrescIdx[i];
        const double val = d.nodePrescVal[i];
        for (int v = 0; v < Q; ++v) ns[v] = (v == idx) ? val : ns[v];
      }
    }
    for (int v = 0; v < Q; ++v) d.nstates[(size_t)v*d.nNodes + n] = ns[v];
  }
};

// ------------------------------------------------------------------------------------------------
// a7: Venkatakrishnan / Barth-Jespersen limiter, evaluated eagerly per cell (the reference evaluates it at
// the first face that touches the cell, FVMCC_PolyRec.cxx:120-167; all its inputs are frozen before the face
// loop, so the value is the same).  Venktn2D.cxx:111-260, Venktn3D.cxx:44-190, BarthJesp.cxx:60-155.
template <int D, class PH = Phys<D> >
struct LimiterK {   // idx over cells
  // 80 registers: measured 8 % faster than the 96 of the default shape; unrolling the loop over the variables at 4 resident
  // blocks (102 registers, all gathers in flight together) was measured 50 % slower (2.88 vs 1.94 ms at 10.6 M cells)
  static constexpr int MINB = 6;
  Dev d; Cfg c;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = PH::Q;
    if (!d.cellActive[cell]) return;
    const bool fresh = (c.residual > c.limitRes) && (c.limitIter > 0 && c.iter < c.limitIter);
    if (!fresh && c.freezeLimiter) return;
    const int fb = d.cellFacePtr[cell], fe = d.cellFacePtr[cell+1];
    const int nbFaces = fe - fb;
    double xc[3] = {0., 0., 0.};
    for (int k = 0; k < D; ++k) xc[k] = d.centres[(size_t)k*d.nStates + cell];
    double dxq[6][3];       // face quadrature point - cell centre
#pragma unroll
    for (int j = 0; j < 6; ++j) if (j < nbFaces) {     // fully unrolled + predicated: dxq / nbrFace stay in registers
      const int f = d.cellFaces[fb + j];
      for (int k = 0; k < D; ++k) dxq[j][k] = d.faceCentres[(size_t)k*d.nFaces + f] - xc[k];
    }
    // neighbour list and average distance: the reference recomputes them inside its loop over the variables
    // (Venktn2D.cxx:150-185); they depend on the geometry only, so they are formed once here, by the same operations
    double avgDistance = 0.0;
    int nbrFace[6] = {0, 0, 0, 0, 0, 0};
    if (c.useFullStencil) {
      const int sb = d.stencilPtr[cell], se = d.stencilPtr[cell+1];
      for (int is = sb; is < se; ++is) {
        const int s = d.stencil[is];
        if (s != cell) {
          double dist = 0.;
          for (int k = 0; k < D; ++k) { const double diff = xc[k] - d.centres[(size_t)k*d.nStates + s]; dist += diff*diff; }
          avgDistance += dsqrt(dist);
        }
      }
      avgDistance /= (se - sb);
    } else {
#pragma unroll
      for (int j = 0; j < 6; ++j) if (j < nbFaces) {
        const int f = d.cellFaces[fb + j];
        const int fl = d.faceL[f];
        const int s = (fl == cell) ? d.faceR[f] : fl;
        nbrFace[j] = s;
        double dist = 0.;
        for (int k = 0; k < D; ++k) { const double diff = xc[k] - d.centres[(size_t)k*d.nStates + s]; dist += diff*diff; }
        avgDistance += dsqrt(dist);
      }
      avgDistance /= nbFaces;
    }
    double weight = 0.0;      // Venktn3DStrict.cxx:87,150-151: from the density (v == 0), kept for the other variables
    for (int v = 0; v < Q; ++v) {
      const double u0 = d.U[(size_t)v*d.nStates + cell];
      double min0 = u0, max0 = u0;
      if (c.useFullStencil) {
        const int sb = d.stencilPtr[cell], se = d.stencilPtr[cell+1];
        for (int is = sb; is < se; ++is) {
          const int s = d.stencil[is];
          if (s != cell) {
            const double us = d.U[(size_t)v*d.nStates + s];
            min0 = dmin(min0, us); max0 = dmax(max0, us);
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 6; ++j) if (j < nbFaces) {
          const double us = d.U[(size_t)v*d.nStates + nbrFace[j]];
          min0 = dmin(min0, us); max0 = dmax(max0, us);
        }
      }
      const double deltaPlusMax = max0 - u0;
      const double deltaPlusMin = min0 - u0;
      double psi = 1.0;       // carried over from face to face when deltaMin == 0 (reference quirk)
      double psimin = 1.1;
      const double ux = d.uX[(size_t)v*d.nCells + cell], uy = d.uY[(size_t)v*d.nCells + cell];
      const double uz = (D == 3) ? d.uZ[(size_t)v*d.nCells + cell] : 0.;
      const bool strict = (c.limiter == CFB200_LIM_VENKTN_STRICT);
      if (strict && v == 0) weight = c.strictCoeff*fabs(ddiv(min0 + 1e-10, max0 + 1e-10)) + (1.0 - c.strictCoeff);
      if ((c.limiter == CFB200_LIM_VENKTN && c.venktnIsMFMHD) || strict) {
        // "implementation from the Venkatakrishnan paper": Venktn3D.cxx:125-156 / Venktn3DStrict.cxx:166-203
        if (!(strict && !c.venktnIsMFMHD)) {
          const double mag = c.magnitudeValues[v];
          const double epsilon2 = pow3(ddiv(c.coeffEps*avgDistance, c.length));
#pragma unroll
          for (int j = 0; j < 6; ++j) if (j < nbFaces) {
            double deltaMin;
            if (D == 2) deltaMin = ux*dxq[j][0] + uy*dxq[j][1];
            else deltaMin = ux*dxq[j][0] + uy*dxq[j][1] + uz*dxq[j][2];
            const bool pos = deltaMin > 0.0;
            // (each quotient with its own reciprocal: sharing one Recip of `mag` across the six unrolled faces gave wrong
            // limiter values on B200 at this launch shape (80 registers, spills) although the expression itself is
            // bit-exact in isolation -- tools/dbg/lim.cu -- and the kernel is right at 2 resident blocks: kept simple)
            const double dMinStar = ((deltaMin < 0.0) ? -1.0 : 1.0)*(ddiv(fabs(deltaMin), mag) + 1e-24);
            const double dPlus = ddiv(pos ? deltaPlusMax : deltaPlusMin, mag);
            const double dPlus2 = dPlus*dPlus;
            const double dPlusMin = dPlus*dMinStar;
            const double Num = (dPlus2 + epsilon2)*dMinStar + ((2*dMinStar)*dMinStar)*dPlus;
            const double Den = ((dPlus2 + (2*dMinStar)*dMinStar) + dPlusMin) + epsilon2;
            const double cand = ddiv(1., dMinStar)*ddiv(Num, Den);
            if (deltaMin > 0.0 || deltaMin < 0.0) psi = cand;        // psi is kept when deltaMin == 0
            if (strict) {
              if (v < 4 || v == 7) psimin = dmin(psi*weight, psimin);
              else psimin = (!c.psiMinEqual1) ? dmin(psi, psimin) : 1.0;
            } else psimin = dmin(psi, psimin);
          }
        }
      } else if (c.limiter == CFB200_LIM_VENKTN) {
        const double epsilon = c.coeffEps*c.magnitudeValues[v]*c.magnitudeValues[v]*pow3(ddiv(avgDistance, c.length));
#pragma unroll
        for (int j = 0; j < 6; ++j) if (j < nbFaces) {
          double deltaMin;
          if (D == 2) deltaMin = ux*dxq[j][0] + uy*dxq[j][1];
          else deltaMin = ux*dxq[j][0] + uy*dxq[j][1] + uz*dxq[j][2];
          // no branches (the divisions of the six faces overlap): numerator and denominator of the candidate that
          // applies are selected first, then divided once: same values as the reference's two if-blocks
          // (Venktn2D.cxx:226-243); psi is kept when deltaMin == 0
          const double dPlusMax2 = deltaPlusMax*deltaPlusMax;
          const double dPlusMaxMin = deltaPlusMax*deltaMin;
          const double numP = dPlusMax2 + 2.0*dPlusMaxMin + epsilon;
          const double denP = dPlusMax2 + dPlusMaxMin + 2.0*deltaMin*deltaMin + epsilon;
          const double y = ddiv(deltaPlusMin, deltaMin);
          const double numM = y*y + 2*y, denM = y*y + y + 2.;
          const bool pos = deltaMin > 0.0;
          const double cand = ddiv(pos ? numP : numM, pos ? denP : denM);
          if (deltaMin > 0.0 || deltaMin < 0.0) psi = cand;
          psimin = dmin(psi, psimin);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 6; ++j) if (j < nbFaces) {
          double deltaMin;
          if (D == 2) deltaMin = ux*dxq[j][0] + uy*dxq[j][1];
          else deltaMin = ux*dxq[j][0] + uy*dxq[j][1] + uz*dxq[j][2];
          const Recip rd = recipOf(deltaMin);
          const double qMax = divBy(deltaPlusMax, rd), qMin = divBy(deltaPlusMin, rd);
          if (deltaMin > 0.0) psi = dmin(1., qMax);
          if (deltaMin < 0.0) psi = dmin(1., qMin);
          psimin = dmin(psi, psimin);
        }
      }
      double* out = &d.lim[(size_t)v*d.nCells + cell];
      // Venktn2D option "2DNoLimiterID" (Venktn2D.cxx:253-258): the limiter of the listed variables is cancelled
      if (D == 2 && c.limiter == CFB200_LIM_VENKTN && c.noLimiterVar[v]) psimin = 1.0;
      *out = fresh ? psimin : dmin(psimin, *out);
    }
  }
};

#include "cfb_face.cuh"
#include "cfb_centred_face.cuh"

// ------------------------------------------------------------------------------------------------
// a21: FVMCC_PseudoSteadyTimeRhs::execute (FVMCC_PseudoSteadyTimeRhs.cxx:106-330):
//   diag = V/dt (dt > 0, unsteady) | updateCoeff/CFL (steady) | V/minDt (useGlobalDT, minDt = min_cells V/updateCoeff*CFL)
//   rhs -= (U - U_past) * diag * resFactor in SOLUTION variables, unless zeroDiagValue[iEq]
//   diagonal block += diag * resFactor * dU/dP, dU/dP numerical (:197-262) or analytical (:266-310)
// The same two device functions serve the stand-alone kernel (cfb200_time_rhs) and the fused form (cfg.fuseTimeRhs:
// CellGather subtracts the rhs part, DiagGather adds the block part), so both give the same bits.
template <int D, class PH>
CFB_HD double timeDiagValue(const Dev& d, const Cfg& c, int cell, double updateCoeff, double minDt) {
  if (c.timeUseGlobalDT) return ddiv(d.volumes[cell], minDt);
  return (c.dt > 0.0) ? ddiv(d.volumes[cell], c.dt) : ddiv(updateCoeff, c.cfl);
}
// dU/dP, analytical: Euler2DPuvtToConsInPuvt.cxx:42-67, Euler3DPvtToConsInPvt.cxx:42-76; identity when update = solution
template <int D, class PH>
CFB_HD double timeAnalyticalEntry(const Cfg& c, const double* u, int i, int j) {
  constexpr int Q = PH::Q;
  if (PH::CENTRED_ONLY || c.updateVar == CFB200_UV_CONS) return (i == j) ? 1. : 0.;
  constexpr int TID = D + 1;
  const double p = u[0] + c.pOffset;
  const double T = u[TID];
  const double rho = ddiv(p, c.R*T);
  const double rhoOvT = ddiv(rho, T);
  const double invRT = ddiv(1., c.R*T);
  double V2;
  if (D == 2) V2 = u[1]*u[1] + u[2]*u[2]; else V2 = u[1]*u[1] + u[2]*u[2] + u[3]*u[3];
  const double gammaMinus1 = c.gamma - 1.;
  double r = 0.;
  if (i == 0) { if (j == 0) r = invRT; else if (j == TID) r = -rhoOvT; }
  else if (i == TID) {
    if (j == 0) r = ddiv(1., gammaMinus1) + 0.5*V2*invRT;
    else if (j == TID) r = -0.5*rhoOvT*V2;
    else r = rho*pick<Q>(u, j);
  } else {
    const double ui = pick<Q>(u, i);
    if (j == 0) r = ui*invRT; else if (j == i) r = rho; else if (j == TID) r = -rhoOvT*ui;
  }
  return r;
}
// what the time term adds to entry (v, iVar) of the diagonal block of a cell with update state u
template <int D, class PH>
CFB_HD double timeBlockEntry(const Cfg& c, const double* u, int v, int iVar, double diagValue) {
  constexpr int Q = PH::Q;
  if (c.timeAnalyticalMatrix) {
    if (c.timeZeroDiag[iVar]) return 0.;        // (sic) the reference tests the flag of the COLUMN in the analytical variant
    return timeAnalyticalEntry<D, PH>(c, u, v, iVar)*diagValue*c.resFactor;
  }
  double up[Q], tempState[Q], tPert[Q];
  const double orig = pick<Q>(u, iVar);
  const double eps = fdEps(c, iVar, orig);
#pragma unroll
  for (int k = 0; k < Q; ++k) up[k] = (k == iVar) ? orig + eps : u[k];
  PH::toCons(c, u, tempState);
  PH::toCons(c, up, tPert);
  const double invEps = ddiv(1., eps);
  double fd = (pick<Q>(tPert, v) - pick<Q>(tempState, v))*invEps;
  fd *= c.timeZeroDiag[v] ? 0.0 : diagValue*c.resFactor;
  return fd;
}
// the whole time-term block of one cell, blk[v*Q + iVar] = value (one thread per cell)
template <int D, class PH>
CFB_HD void timeBlock(const Cfg& c, double* u, double diagValue, double* blk, int stride) {
  constexpr int Q = PH::Q;
  if (c.timeAnalyticalMatrix) {
    for (int iVar = 0; iVar < Q; ++iVar)
      for (int v = 0; v < Q; ++v) blk[(size_t)(v*Q + iVar)*stride] = timeBlockEntry<D, PH>(c, u, v, iVar, diagValue);
    return;
  }
  double tempState[Q], tPert[Q];
  PH::toCons(c, u, tempState);
  for (int iVar = 0; iVar < Q; ++iVar) {
    const double orig = u[iVar];
    const double eps = fdEps(c, iVar, orig);
    u[iVar] = orig + eps;
    PH::toCons(c, u, tPert);
    const double invEps = ddiv(1., eps);
    for (int v = 0; v < Q; ++v) {
      double fd = (tPert[v] - tempState[v])*invEps;
      fd *= c.timeZeroDiag[v] ? 0.0 : diagValue*c.resFactor;
      blk[(size_t)(v*Q + iVar)*stride] = fd;
    }
    u[iVar] = orig;
  }
}
// rhs part for one cell (AoS rhs row r[] in registers)
template <int D, class PH>
CFB_HD void timeRhsRow(const Dev& d, const Cfg& c, int cell, const double* u, double diagValue, double* r) {
  constexpr int Q = PH::Q;
  double up[Q], tempState[Q], tPast[Q];
#pragma unroll
  for (int v = 0; v < Q; ++v) up[v] = d.pastU[(size_t)v*d.nCells + cell];
  PH::toCons(c, u, tempState);
  PH::toCons(c, up, tPast);
#pragma unroll
  for (int v = 0; v < Q; ++v) {
    const double dU = tempState[v] - tPast[v];
    r[v] -= c.timeZeroDiag[v] ? 0.0 : dU*diagValue*c.resFactor;
  }
}


// Data.SourceTerm = MHDConsACAST (FiniteVolumeMHD/MHDConsACASourceTerm.cxx:164-211, solar terms off): the damping term of the
// divergence-cleaning potential, from the CELL state u; correctionType "Hyperbolic" is mhdDissipCoeff = 0
template <int D, class PH>
CFB_HD void sourceTerm(const Dev& d, const Cfg& c, int cell, const double* u, double* S) {
  constexpr int Q = PH::Q;
  for (int v = 0; v < Q - 1; ++v) S[v] = 0.;
  double last = 0.;
  if (c.mhdDissipCoeff != 0.0) {
    const double refSpeedSq = c.refSpeed*c.refSpeed;
    const double dissipCoeffSq = c.mhdDissipCoeff*c.mhdDissipCoeff;
    last = (-ddiv(refSpeedSq, dissipCoeffSq))*u[Q-1]*d.volumes[cell];
  }
  S[Q-1] = last;
}
// entry (row, col) of the finite-difference Jacobian of the source on the diagonal block (FVMCC_ComputeRhsJacob.cxx:359-364
// with addSourceTermNumJacob, .hh:96-105; _upStFactor = -resFactor, FVMCC_ComputeRHS.hh:143)
template <int D, class PH>
CFB_HD double sourceJacobEntry(const Dev& d, const Cfg& c, int cell, int row, int col) {
  constexpr int Q = PH::Q;
  double u[Q], S[Q], Sp[Q];
  for (int v = 0; v < Q; ++v) u[v] = d.U[(size_t)v*d.nStates + cell];
  sourceTerm<D, PH>(d, c, cell, u, S);
  double orig = 0.;
  for (int v = 0; v < Q; ++v) orig = (v == col) ? u[v] : orig;
  const double eps = fdEps(c, col, orig);
  for (int v = 0; v < Q; ++v) u[v] = (v == col) ? u[v] + eps : u[v];
  sourceTerm<D, PH>(d, c, cell, u, Sp);
  const double invEps = ddiv(1., eps);
  double sb = 0., sp = 0.;
  for (int v = 0; v < Q; ++v) { sb = (v == row) ? S[v] : sb; sp = (v == row) ? Sp[v] : sp; }
  double sum = 0.;
  sum += (sp - sb)*invEps;
  return sum*(-c.resFactor);
}

// a17: FVMCC_ComputeRHS::updateRHS (FVMCC_ComputeRHS.cxx:455-514) as a gather: for each cell, its faces in
// ascending loop order; rhs starts from 0, updateCoeff accumulates onto the incoming value
// (it is zeroed by the updater, UpdateSol.cxx:192).
// FUSED: also applies the pseudo-steady time term (cfg.fuseTimeRhs); a template parameter so that the plain gather
// keeps its register count (56 -> occupancy) when the term runs as its own kernel
// SRC: Data.SourceTerm, added (times resFactor) when the face loop first touches the cell (FVMCC_ComputeRHS.cxx:538-600)
template <int D, class PH = Phys<D>, bool FUSED = false, bool SRC = false>
struct CellGather {   // idx over cells
  static constexpr int BLK = PH::Q*PH::Q;      // FUSED: doubles of per-thread staging (the time-term block of the cell)
  Dev d; Cfg c;
  CFB_HD double* out() const { return d.bsrVal; }
  CFB_HD int slot(int cell) const { return (c.computeJacobian && d.updatable[cell]) ? d.diagSlot[cell] : -1; }
  CFB_HD void operator()(int cell) const { run(cell, nullptr, 0); }
  // stg (entry k at stg[k*stride]): where the FUSED variant leaves the time-term block of the cell for the block-wide
  // coalesced write-out of k_cell_staged (cfb_capi.cu) into the cell's diagonal slot
  CFB_HD void run(int cell, double* stg, int stride) const {
    constexpr int Q = PH::Q;
    double r[Q];
    for (int v = 0; v < Q; ++v) r[v] = 0.;
    const int rc = refId(d, cell);
    double uc = d.updateCoeff[rc];
    const double coeff = c.resFactor*1.0;
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    bool first = true; (void)first;
    for (int j = d.cellFacePtr[cell]; j < d.cellFacePtr[cell+1]; ++j) {
      const int e = d.cellFaceSorted[j];
      const bool isR = e < 0;
      const int f = isR ? ~e : e;
      const int oL = d.faceL[f];
      const bool skip = !(d.updatable[oL] || (f < d.nInner && d.updatable[d.faceR[f]]));
      if (skip) continue;
      if constexpr (SRC) {
        if (first) {
          first = false;
          double u[Q], S[Q];
          for (int v = 0; v < Q; ++v) u[v] = d.U[(size_t)v*d.nStates + cell];
          sourceTerm<D, PH>(d, c, cell, u, S);
          for (int v = 0; v < Q; ++v) r[v] += (c.resFactor*S[v])*1.0;
        }
      }
      for (int v = 0; v < Q; ++v) {
        const double rFlux = coeff*d.faceFlux[(size_t)v*d.nFaces + f];
        if (isR) r[v] += rFlux*1.0; else r[v] -= rFlux*1.0;
      }
      uc += isR ? d.ucR[f] : d.ucL[f];
      if (hasDiff) uc += d.ucD[f];
    }
    if (FUSED && d.updatable[cell]) {     // a21 fused: FVMCC_PseudoSteadyTimeRhs on the finished residual row
      double u[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) u[v] = d.U[(size_t)v*d.nStates + cell];
      const double diagValue = timeDiagValue<D, PH>(d, c, cell, uc, 0.);
      timeRhsRow<D, PH>(d, c, cell, u, diagValue, r);
      // block part -> staging area -> (coalesced) the not yet written diagonal slot: DiagGather adds it to its face sum
      if (c.computeJacobian) timeBlock<D, PH>(c, u, diagValue, stg, stride);
    }
    for (int v = 0; v < Q; ++v) d.rhs[(size_t)rc*Q + v] = r[v];
    d.updateCoeff[rc] = uc;
  }
};

// a22: FVMCC_ComputeRHSMHD<BASE>::computeInterConvDiff (FiniteVolumeMHD/FVMCC_ComputeRHSMHD.ci:160-203): the cell divergence of B
// from the last component (the phi equation) of the face fluxes of the last evaluation,
//   divB[L] += F[q-1]/(refSpeed^2 V_L),  divB[R] -= F[q-1]/(refSpeed^2 V_R),
// gathered per cell in face-loop order like the residual.
template <int D, class PH = Phys<D> >
struct DivBGather {   // idx over cells
  Dev d; Cfg c; double* out;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = PH::Q;
    const double den = (c.refSpeed*c.refSpeed)*d.volumes[cell];
    const Recip rd = recipOf(den);
    double acc = 0.;
    for (int j = d.cellFacePtr[cell]; j < d.cellFacePtr[cell+1]; ++j) {
      const int e = d.cellFaceSorted[j];
      const bool isR = e < 0;
      const int f = isR ? ~e : e;
      const bool skip = !(d.updatable[d.faceL[f]] || (f < d.nInner && d.updatable[d.faceR[f]]));
      if (skip) continue;
      const double t = divBy(d.faceFlux[(size_t)(Q-1)*d.nFaces + f], rd);
      if (isR) acc -= t; else acc += t;
    }
    out[refId(d, cell)] = acc;
  }
};

// Diagonal blocks: one thread per (cell, block entry).  diag(c) = sum over the cell's faces in loop order of
//   inner face, c == L:  -block(R,c)      inner face, c == R:  -block(L,c)      boundary face: +bnd block
// Rows of non-updatable cells are staging only and are zeroed by ZeroRows afterwards.
// SRC: the finite-difference Jacobian of Data.SourceTerm lands on the block right after the cell's first face
template <int D, class PH = Phys<D>, bool FUSED = false, bool SRC = false>
struct DiagGather {   // idx over nCells*GS: one thread per (cell, block entry), no idle lanes
  static constexpr int GS = PH::Q*PH::Q;
  // 16 resident 128-thread blocks = 2048 threads per SM = 32 registers: measured 4.15 ms at 90 % occupancy against 4.95 ms
  // at the 36 registers (67 %) the unconstrained allocation sometimes picks (220^3, fused time term)
  static constexpr int MINB = 16;
  Dev d; int cell0;      // cell0: first cell of this launch (chunks of < 2^31 entries start at a cell boundary)
  Cfg c;                 // read by the SRC variant only
  CFB_HD void operator()(int idx) const {
    constexpr int Q = PH::Q;
    const int lc = idx/GS, e = idx - lc*GS;
    const int cell = cell0 + lc;
    if (e >= Q*Q) return;
    double acc = 0.;
    if (d.updatable[cell]) {
      // diagSrc[j] (built once on the host, aligned with cellFaceSorted): the BSR slot of the block this face
      // contributes with a minus sign, or ~g for boundary face g (plus sign).  Batches of 6 faces: all source
      // blocks of a batch are requested before the first is consumed (memory-level parallelism), then added in
      // face-loop order.
      const int jb = d.cellFacePtr[cell], je = d.cellFacePtr[cell+1];
      for (int j0 = jb; j0 < je; j0 += 6) {
        double x[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          x[k] = 0.;
          if (j0 + k < je) {
            const int s = d.diagSrc[j0 + k];
            x[k] = (s >= 0) ? d.bsrVal[(size_t)s*Q*Q + e] : d.bndBlocks[(size_t)(~s)*Q*Q + e];
          }
        }
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (j0 + k < je) {
            if (d.diagSrc[j0 + k] >= 0) acc -= x[k]; else acc += x[k];
            if constexpr (SRC) { if (j0 + k == jb) acc += sourceJacobEntry<D, PH>(d, c, cell, e/Q, e - (e/Q)*Q); }
          }
      }
      // a21 fused: CellGather (one thread per cell, updateCoeff complete there) left the time-term block in the diagonal
      // slot; it lands on the finished face sum, as FVMCC_PseudoSteadyTimeRhs adds it after the face loop
      if (FUSED) acc += d.bsrVal[(size_t)d.diagSlot[cell]*Q*Q + e];
    }
    d.bsrVal[(size_t)d.diagSlot[cell]*Q*Q + e] = acc;
  }
};

template <int D, class PH = Phys<D> >
struct ZeroRows {   // idx over (non-updatable cell list) * GS
  static constexpr int GS = PH::Q*PH::Q;
  Dev d; const int* cells; int first;      // first: position in the list of the first cell of this launch
  CFB_HD void operator()(int idx) const {
    constexpr int Q = PH::Q;
    const int row = refId(d, cells[first + idx/GS]), e = idx - (idx/GS)*GS;
    if (e >= Q*Q) return;
    for (int k = d.bsrRowPtr[row]; k < d.bsrRowPtr[row+1]; ++k) d.bsrVal[(size_t)k*Q*Q + e] = 0.;
  }
};

// ------------------------------------------------------------------------------------------------
// a21 kernels (the device functions are above CellGather)
template <int D, class PH = Phys<D> >
struct TimeRhs {   // idx over cells
  // q = 4, 5: 106-126 registers without spills, 16 resident warps per SM; q = 9 is faster at the default shape (measured)
  static constexpr int MINB = (PH::Q > 5) ? 2 : 4;
  Dev d; Cfg c; int doJac; double minDt;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = PH::Q;
    if (!d.updatable[cell]) return;
    const int rc = refId(d, cell);
    const double diagValue = timeDiagValue<D, PH>(d, c, cell, d.updateCoeff[rc], minDt);
    double u[Q], r[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { u[v] = d.U[(size_t)v*d.nStates + cell]; r[v] = d.rhs[(size_t)rc*Q + v]; }
    timeRhsRow<D, PH>(d, c, cell, u, diagValue, r);
#pragma unroll
    for (int v = 0; v < Q; ++v) d.rhs[(size_t)rc*Q + v] = r[v];
    if (!doJac) return;
    double* blk = &d.bsrVal[(size_t)d.diagSlot[cell]*Q*Q];
    if (c.timeAnalyticalMatrix) {
      for (int iVar = 0; iVar < Q; ++iVar)
        for (int v = 0; v < Q; ++v) blk[v*Q + iVar] += timeBlockEntry<D, PH>(c, u, v, iVar, diagValue);
      return;
    }
    double tempState[Q], tPert[Q];
    PH::toCons(c, u, tempState);
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = u[iVar];
      const double eps = fdEps(c, iVar, orig);
      u[iVar] = orig + eps;
      PH::toCons(c, u, tPert);
      const double invEps = ddiv(1., eps);
      for (int v = 0; v < Q; ++v) {
        double fd = (tPert[v] - tempState[v])*invEps;
        fd *= c.timeZeroDiag[v] ? 0.0 : diagValue*c.resFactor;
        blk[v*Q + iVar] += fd;
      }
      u[iVar] = orig;
    }
  }
};
// useGlobalDT: per-cell V/updateCoeff*CFL (the reference's minimum runs over every local state, :121-127); the minimum of
// the values is exact whatever the reduction order
template <int D, class PH = Phys<D> >
struct TimeLocalDt {   // idx over cells
  Dev d; Cfg c;
  CFB_HD void operator()(int cell) const { d.partial[cell] = ddiv(d.volumes[cell], d.updateCoeff[refId(d, cell)])*c.cfl; }
};

// ------------------------------------------------------------------------------------------------
// §8f-1: ForwardEuler UpdateSol.cxx:83-193 on the device: dt = CFL/updateCoeff, U += rhs*dt (rhs itself is left unscaled),
// updateCoeff = 0; per-cell squared dU[monitorVar] goes to `partial` for the ordered norm reduction.
template <int D, class PH = Phys<D> >
struct UpdateSolFwdEuler {   // idx over cells
  Dev d; double cfl; int monitorVar;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = PH::Q;
    double sq = 0.;
    const int rc = refId(d, cell);
    if (!d.updatable[cell]) {
      for (int v = 0; v < Q; ++v) d.rhs[(size_t)rc*Q + v] = 0.;
    } else {
      const double uc = d.updateCoeff[rc];
      const bool isZero = fabs(uc) <= 2.2250738585072014e-308;
      const double dt = isZero ? 0. : cfl/uc;
      for (int v = 0; v < Q; ++v) {
        double rhsState = d.rhs[(size_t)rc*Q + v];     // a copy: the rhs socket itself is NOT scaled (UpdateSol.cxx:176-177)
        d.U[(size_t)v*d.nStates + cell] += rhsState *= dt;
      }
      const double t = d.rhs[(size_t)rc*Q + monitorVar]; // ComputeL2Norm.cxx:111-125 on the rhs socket
      sq = t*t;
    }
    d.updateCoeff[rc] = 0.0;
    d.partial[rc] = sq;              // the ordered norm sum runs in reference order
  }
};

// §8f-1: NewtonMethod StdUpdateSol.cxx:110-141,256 on the device: the solution increment dU of the linear solve sits in
// the rhs socket; U += alpha[iEq]*dU for the updatable states, dU = 0 for the others, updateCoeff = 0; per-cell squared
// dU[monitorVar] goes to `partial` for the ordered norm reduction (ComputeL2Norm.cxx:111-125 on the rhs socket).
template <int D, class PH = Phys<D> >
struct UpdateSolNewton {   // idx over cells
  Dev d; double alpha[PH::Q]; int monitorVar;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = PH::Q;
    double sq = 0.;
    const int rc = refId(d, cell);
    if (!d.updatable[cell]) {
      for (int v = 0; v < Q; ++v) d.rhs[(size_t)rc*Q + v] = 0.;
    } else {
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        const double dU = d.rhs[(size_t)rc*Q + v];
        d.U[(size_t)v*d.nStates + cell] += alpha[v]*dU;
        if (v == monitorVar) sq = dU*dU;
      }
    }
    d.updateCoeff[rc] = 0.0;
    d.partial[rc] = sq;
  }
};

} // namespace cfb
