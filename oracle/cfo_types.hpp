// ORACLE — TEST INFRASTRUCTURE ONLY.
// CPU restatement of COOLFluiD's cell-centred finite-volume (FVMCC) ComputeRHS /
// ComputeRhsJacob path.  Nothing under oracle/ is part of the shipped product:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load it, and only as the checker / CPU baseline.
//
// Parity status: the arithmetic follows the reference file:line cited at each
// function; the evaluation order of the lazy MathTools expression templates is
// pinned against the reference's own headers through oracle/_ref (see
// oracle/ref_mathtools.cpp); the end-to-end behaviour is pinned against the
// reference's regression value `### Residual = -4.0017502` of
// plugins/NavierStokes/testcases/Wedge/wedgeFVM.CFcase:14 (tests/test_wedge_pin.py).
// Per-cell residual / Jacobian entries have no golden vectors in the reference
//  and are pinned only through the above.
//
// All paths are relative to /root/reference.
#pragma once
#include <cstdint>
#include <cmath>
#include <vector>
#include <array>
#include <algorithm>
#include <stdexcept>
#include <string>
#include <limits>

namespace cfo {

constexpr int MAXEQ = 9;
constexpr int MAXDIM = 3;

// ---- enums shared (by value) with the product's include/cfb200.h -------------
enum PhysicalModel { PM_EULER = 0, PM_MHD3D_PROJECTION = 1 };
enum UpdateVar { UV_CONS = 0, UV_PVT = 1 };            // Euler{2,3}DCons / Euler2DPuvt, Euler3DPvt
enum FluxSplitter { FS_ROE = 0, FS_AUSMPLUS = 1, FS_AUSMPLUSUP = 2, FS_LAXFRIED = 3, FS_HLL = 4, FS_ROESA = 5, FS_CENTRED = 6 };
enum PolyRec { PR_CONSTANT = 0, PR_LINEARLS = 1 };
enum Limiter { LIM_NONE = 0, LIM_VENKTN = 1, LIM_BARTHJESP = 2, LIM_VENKTN_STRICT = 3 };
enum Diffusive { DIFF_NONE = 0, DIFF_NS = 1 };
enum ViscLaw { VISC_SIMPLER_SUTHERLAND = 0, VISC_SUTHERLAND = 1, VISC_FIXED = 2 };
enum BCType {
  BC_SUPER_OUTLET = 0,     // FiniteVolume/SuperOutlet.cxx:38-47        ghost = inner
  BC_DIRICHLET = 1,        // ghost[v] = 2*value[v] - inner[v] for v in mask, copy otherwise
                           //   SuperInlet.cxx:156-209, SubInletEulerPvtVT.ci:52-77,
                           //   SubOutletEulerPvt.ci:36-51, NoSlipWallAdiabaticNSPvt.cxx:109-128
  BC_MIRROR_VELOCITY = 2,  // FiniteVolume/MirrorVelocity.cxx:71-111
  BC_MIRROR_EULER_CONS = 3, // FiniteVolumeNavierStokes/MirrorEuler2D.cxx:56-94 (+3D analogue)
  BC_FARFIELD_EULER = 4,   // FiniteVolumeNavierStokes/FarFieldEuler2D.cxx:127-300, FarFieldEuler3D.cxx:62-196; value = p u v [w] T at infinity
  BC_FARFIELD_EULER_CHAR = 7, // FiniteVolumeNavierStokes/FarFieldEulerChar2D.cxx:36-254, FarFieldEulerChar3D.cxx:40-200; value as BC_FARFIELD_EULER
  BC_SUPER_OUTLET_MHD = 6, // FiniteVolumeMHD/SuperOutletMHD3DProjection.cxx:76-239; mask[0] = coronalBC, value[0] = refPhi
  BC_NOSLIP_ISOTHERMAL = 5 // FiniteVolumeNavierStokes/NoSlipWallIsothermalNSvt.ci:178-243; value = TGhostMin, uWall[D], TWall (adimensional)
};

// One boundary condition (one per boundary TRS).
struct BC {
  int32_t type;
  int32_t mask[MAXEQ];        // BC_DIRICHLET: which variables are imposed
  double value[MAXEQ];        // BC_DIRICHLET: imposed (adimensional, update-variable) values
  int32_t zeroGrad[MAXEQ];    // ZeroGradientFlags (FVMCC_ComputeRHS.cxx:186-188)
};

// Run configuration: everything the reference reads from the .CFcase + SubSystemStatus.
struct Config {
  int32_t dim;            // 2 | 3
  int32_t nEqs;           // dim + 2
  int32_t updateVar;      // UpdateVar
  int32_t fluxSplitter;   // FluxSplitter
  int32_t polyRec;        // PolyRec
  int32_t limiter;        // Limiter
  int32_t diffusive;      // Diffusive
  int32_t viscLaw;        // ViscLaw
  int32_t computeJacobian;// 0 | 1 (numerical, FVMCC_ComputeRhsJacob) | 2 ("NumJacobAnalytic": FVMCC_ComputeRhsJacobAnalytic)
  int32_t useFullStencil; // limiter option
  int32_t freezeLimiter;  // PolyReconstructor.ci:66
  int32_t stopLimiting;   // FVMCC_PolyRec.cxx:52
  int32_t isAdimensional; // NavierStokesPhysicalModel.ci:137
  int32_t freezeDiffCoeff;// FVMCC_ComputeRhsJacob option
  int32_t iter;           // SubSystemStatus::getNbIter()
  int32_t limitIter;      // PolyReconstructor.ci:57
  double gamma;           // EulerTerm.cxx:48
  double R;               // EulerTerm::getR()  (_RRef)
  double pOffset;         // _p0Inf/m_refPhysicalData[P]  (EulerTerm.hh:199-202)
  double refValues[MAXEQ];// PhysicalModel refValues -> NumericalJacobian (RoeFlux.cxx:113-115)
  double fdTol;           // NumericalJacobian.cxx:31  (10e-7)
  double resFactor;       // FVMCC_ComputeRHS.hh:98
  double residual;        // SubSystemStatus::getResidual()
  double limitRes;        // PolyReconstructor.ci:60
  double gradientFactor;  // PolyReconstructor.ci:63
  double coeffEps;        // Venktn2D.cxx:58
  double length;          // Venktn2D.cxx:62 / refLength
  double magnitudeValues[MAXEQ]; // Venktn2D.cxx:96-106
  double diffRedCoeff;    // RoeFlux.cxx:62 DiffCoeff
  // Navier-Stokes terms (NavierStokesPhysicalModel.ci:69-170, NSTerm)
  double coeffTau, coeffQ, cpOverPrandtl, muRef /*refDataDiff[MU]*/, tempRef, pRef, artDiffCoeff;
  double viscRef, viscTRef, viscSuthConst, fixedVisc;
  double cfl;             // for the pseudo-steady time term
  // AUSM family (FiniteVolumeNavierStokes/AUSMFlux.ci:55, AUSMPlusFlux.ci:17-23, AUSMPlusUpFlux.ci:20-33)
  int32_t choiceA12;      // interface sound speed: 1 Liou critical, 2 arithmetic mean, 3 geometric mean
  int32_t hllUseAlpha;    // HLLFlux.cxx:48 UseAlpha
  double ausmBeta, ausmAlpha;                 // AUSM+ : 1/8, 3/16
  double ausmKu, ausmKp, ausmSigma, machInf;  // AUSM+up : 0.75, 0.25, 1.0, (required > 0)
  // physical model: 0 = Euler / Navier-Stokes (nEqs = dim + 2), 1 = MHD3DProjection (dim 3, nEqs 9, Cons variables)
  int32_t physicalModel;
  int32_t entropyFixID;   // RoeEntropyFixFlux.cxx:35 (0 = plain RoeFlux, 1 | 2 = "RoeEntropyFix")
  double refSpeed;        // MHDProjectionTerm refSpeed (hyperbolic divergence cleaning speed)
  double coeffSigmoid;    // AUSMFlux.ci:64 coeffSigmoid (1e-8): blending of the sigmoid switches of the analytic Jacobian
  int32_t derivativeStrategy;  // 0 DiamondVolume{2,3}D (CellCenterFVMData.cxx:298 default), 1 Corrected{2,3}D, 2 StateDiff, 3 CorrectedGG{2,3}D
  int32_t fuseTimeRhs;         // product-side scheduling flag (no meaning for the oracle)
  // FVMCC_StdComputeTimeRhs.cxx:42-57 zeroDiagValue; FVMCC_PseudoSteadyTimeRhs.cxx:40-46 useAnalyticalMatrix, useGlobalDT
  int32_t timeZeroDiag[MAXEQ];
  int32_t timeAnalyticalMatrix;
  int32_t timeUseGlobalDT;
  int32_t renumber;            // product-side storage permutation (no meaning for the oracle)
  double dt;                   // SubSystemStatus::getDT()
  int32_t venktnIsMFMHD;       // Venktn2D/3D option isMFMHD (Venktn3D.cxx:125-156)
  int32_t psiMinEqual1;        // Venktn3DStrict.cxx:33,47
  double strictCoeff;          // Venktn3DStrict.cxx:31,44
  int32_t roeSaOldStencil;     // RoeSAFlux.cxx:44-45
  int32_t mhdPotentialBType;   // MHDTerm.cxx:25: 0 none, 1 "Dipole"
  double mhdDipole[3];         // MHDTerm.cxx:38-40 mX mY mZ
  int32_t noLimiterVar[MAXEQ]; // Venktn2D.cxx:40-41,253-258 "2DNoLimiterID" as a per-variable flag
  int32_t correctedGGUseWeights; // CorrectedDerivativeGG2D.cxx:60-61
  int32_t sourceTerm;          // Data.SourceTerm: 0 none, 1 "MHDConsACAST" (solar terms off)
  double mhdDissipCoeff;       // MHDTerm dissipCoeff when correctionType == "Mixed", 0 for "Hyperbolic"
};

struct MeshView {
  int32_t dim, nCells, nGhosts, nNodes, nFaces, nInner;
  const double* nodes;      // [nNodes*dim]
  const double* centres;    // [(nCells+nGhosts)*dim]: cell centres, then ghost centres
  const int32_t* faceL;     // [nFaces]  faces are stored in the reference's LOOP order:
  const int32_t* faceR;     //           InnerFaces first, then each boundary TRS; R = nCells+ghost for boundary faces
  const int32_t* faceNodePtr; // [nFaces+1]
  const int32_t* faceNodes;
  const double* normals;    // [nFaces*dim]   area-weighted, outward from L
  const double* areas;      // [nFaces]
  const double* volumes;    // [nCells]
  const int32_t* cellFacePtr; // [nCells+1]
  const int32_t* cellFaces; // loop-order face index, in the cell's local face order
  const int32_t* stencilPtr;  // [nCells+1]
  const int32_t* stencil;   // state indices (cells, or nCells+ghost)
  const int32_t* nodeNbrPtr;  // [nNodes+1]
  const int32_t* nodeNbr;   // state indices
  const double* nodeW;      // nodal extrapolation weights (DistanceBasedExtrapolator.ci:105-125)
  const double* cvol;       // [nFaces] diamond control volumes (DiamondVolume*Derivative)
  const uint8_t* cvSign;    // [nFaces] _changeNormalSign
  const int32_t* faceBC;    // [nFaces] -1 inner, else index into the BC array
  const uint8_t* updatable; // [nCells] isParUpdatable
  const int32_t* cellNodePtr; // [nCells+1] cell -> nodes in the cell-local order (CorrectedGG only; may be NULL otherwise)
  const int32_t* cellNodes;
  // position-dependent Dirichlet data (SuperInlet.cxx:156-209 with a parsed Def function evaluated at the boundary
  // point of each face): when bcFaceFlag[ghost] is set, BC_DIRICHLET reads bcFaceVal[ghost*q + v] instead of BC::value
  const double* bcFaceVal;    // [nGhosts*q] or NULL
  const uint8_t* bcFaceFlag;  // [nGhosts] or NULL
  // NodalExtrapolation = DistanceBasedGMove (DistanceBasedExtrapolatorGMove.cxx): _isNodeToPrescribe, ValuesIdx, Values
  const uint8_t* nodePresc;   // [nNodes] or NULL
  const int32_t* nodePrescIdx;  // [nNodePresc]
  const double* nodePrescVal;   // [nNodePresc]
  int32_t nNodePresc;
};

} // namespace cfo
