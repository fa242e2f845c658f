/* cfb200 — C ABI of the B200-native FVMCC ComputeRHS / ComputeRhsJacob engine.
 *
 * This header is the drop-in boundary: a thin COOLFluiD plugin
 * (plugin/FiniteVolumeB200, a MethodCommand<CellCenterFVMData> registered with
 * MethodCommandProvider exactly like plugins/FiniteVolumeCUDA/FVMCC_ComputeRHSCell.hh:31)
 * copies socket data in, calls these entry points, and copies rhs / updateCoeff /
 * the block Jacobian out.  Plain pointers and sizes only; no exceptions cross it
 * (0 = success, non-zero = error code, text through cfb200_last_error).
 *
 * Every entry point cites the reference interface it replaces (paths relative to
 * the COOLFluiD tree).  All reals are FP64 (CFreal = double, src/Common/COOLFluiD.hh:92-106),
 * all indices 32-bit (CFuint without CF_HAVE_LONG).
 *
 * The library needs an NVIDIA B200 (sm_100a).  There is NO CPU fallback: every
 * compute entry point returns CFB200_ERR_NO_DEVICE when no CUDA device is usable.
 */
#ifndef CFB200_H
#define CFB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CFB200_MAXEQ 9
#define CFB200_ABI_VERSION 3

/* ---- enums: values are shared with the .CFcase option strings through the plugin ---- */
enum cfb200_source_term { CFB200_ST_NONE = 0, CFB200_ST_MHD_CONS_ACA = 1 };
enum cfb200_physical_model { CFB200_PM_EULER = 0, CFB200_PM_MHD3D_PROJECTION = 1 };
enum cfb200_update_var { CFB200_UV_CONS = 0 /* Euler2DCons, Euler3DCons */, CFB200_UV_PVT = 1 /* Euler2DPuvt, Euler3DPvt */ };
enum cfb200_flux_splitter { CFB200_FS_ROE = 0 /* "Roe" plugins/FiniteVolume/RoeFlux.cxx */,
                            CFB200_FS_AUSMPLUS = 1 /* "AUSMPlus2D/3D" FiniteVolumeNavierStokes/AUSMPlusFlux.ci */,
                            CFB200_FS_AUSMPLUSUP = 2 /* "AUSMPlusUp2D/3D" AUSMPlusUpFlux.ci:50-135 */,
                            CFB200_FS_LAXFRIED = 3 /* "LaxFried" plugins/FiniteVolume/LaxFriedFlux.cxx:79-190 */,
                            CFB200_FS_HLL = 4 /* "HLL" plugins/FiniteVolume/HLLFlux.cxx:73-150 */,
                            CFB200_FS_ROESA = 5 /* "RoeSA" plugins/FiniteVolume/RoeSAFlux.cxx:66-197: Roe with Sanders'
                                                   H-correction; entropyFixID then is RoeSA's own option (0 | 3: |ev| + eta,
                                                   1: Harten-like with eta, 2: max(|ev|, eta)) */,
                            CFB200_FS_CENTRED = 6 /* "Centred" plugins/FiniteVolume/CentredFlux.cxx:55-90: 0.5*(fR + fL), no
                                                     dissipation (Euler variable sets and MHD3DProjection) */ };
enum cfb200_poly_rec { CFB200_PR_CONSTANT = 0 /* "Constant" ConstantPolyRec.cxx */,
                       CFB200_PR_LINEARLS = 1 /* "LinearLS2D"/"LinearLS3D" LeastSquareP1PolyRec{2,3}D.cxx */ };
enum cfb200_limiter { CFB200_LIM_NONE = 0, CFB200_LIM_VENKTN = 1 /* "Venktn2D"/"Venktn3D" */,
                      CFB200_LIM_BARTHJESP = 2 /* "BarthJesp" */, CFB200_LIM_VENKTN_STRICT = 3 /* "Venktn3DStrict" */ };
enum cfb200_diffusive { CFB200_DIFF_NONE = 0, CFB200_DIFF_NS = 1 /* "NavierStokes" NSFlux.ci + DiamondVolume{2,3}D */ };
enum cfb200_visc_law { CFB200_VISC_SIMPLER_SUTHERLAND = 0, CFB200_VISC_SUTHERLAND = 1, CFB200_VISC_FIXED = 2 };
enum cfb200_bc_type {
  CFB200_BC_SUPER_OUTLET = 0,      /* FiniteVolume/SuperOutlet.cxx:38-47 */
  CFB200_BC_DIRICHLET = 1,         /* ghost[v] = 2*value[v]-inner[v] where mask[v]: SuperInlet.cxx:156-209 (constant Def),
                                      SubInletEulerPvtVT.ci:52-77, SubOutletEulerPvt.ci:36-51, NoSlipWallAdiabaticNSPvt.cxx:109-128 */
  CFB200_BC_MIRROR_VELOCITY = 2,   /* FiniteVolume/MirrorVelocity.cxx:71-111 */
  CFB200_BC_MIRROR_EULER_CONS = 3, /* FiniteVolumeNavierStokes/MirrorEuler2D.cxx:56-94, MirrorEuler3D */
  CFB200_BC_FARFIELD_EULER = 4,    /* FiniteVolumeNavierStokes/FarFieldEuler2D.cxx:127-300 (conservative variables),
                                      FarFieldEuler3D.cxx:62-196 (Cons or Pvt): value = p, u, v, [w,] T at infinity */
  CFB200_BC_FARFIELD_EULER_CHAR = 7, /* FiniteVolumeNavierStokes/FarFieldEulerChar2D.cxx:36-254, FarFieldEulerChar3D.cxx:40-200 (Cons, Puvt,
                                      Pvt; no StateInterpolator): value as CFB200_BC_FARFIELD_EULER; the subsonic branches follow
                                      the mean of the two Riemann invariants */
  CFB200_BC_SUPER_OUTLET_MHD = 6,  /* FiniteVolumeMHD/SuperOutletMHD3DProjection.cxx:76-239 (MHD3DProjection only; 16 of the reference's
                                      cases): mask[0] = option coronalBC (default true: r^2 Br, Btheta, r Bphi, r^3 rho, r^2 rho Vr,
                                      rho Vtheta, r Vphi, r^3 p continuous over the inner / ghost state positions, phi mirrored),
                                      else the inner physical data with phi = value[0] (option refPhi) */
  CFB200_BC_NOSLIP_ISOTHERMAL = 5  /* FiniteVolumeNavierStokes/NoSlipWallIsothermalNSvt.ci:178-243 (Puvt / Pvt variables): velocity and
                                      temperature linearly extrapolated through the wall values over the inner-wall / inner-ghost
                                      distances of FVMCC_BaseBC.ci:271-350, the other variables copied.  value[1..dim] = uWall,
                                      value[dim+1] = TWall / tempRef, value[0] = TGhostMin / tempRef.  When the ghost temperature
                                      2 TWall - T_inner falls below TGhostMin the reference MOVES the ghost node (repositionNode,
                                      FVMCC_BaseBC.hh:290-340): not done here, the evaluation reports CFB200_ERR_GHOST_REPOSITION */
};

enum cfb200_error {
  CFB200_OK = 0,
  CFB200_ERR_BAD_VALUE = 1,   /* a^2 < 0 in the Roe linearisation: the reference throws BadValueException (Euler2DLinearRoe.cxx:79-85) */
  CFB200_ERR_INVALID_ARG = 2,
  CFB200_ERR_NO_DEVICE = 3,   /* no usable CUDA device: the product has no CPU path */
  CFB200_ERR_CUDA = 4,
  CFB200_ERR_STATE = 5,       /* call order (mesh/states not set) */
  CFB200_ERR_UNSUPPORTED = 6,
  CFB200_ERR_NCCL = 7,
  CFB200_ERR_GHOST_REPOSITION = 8  /* an isothermal wall would reposition a ghost node (see CFB200_BC_NOSLIP_ISOTHERMAL) */
};

/* One boundary condition per boundary TRS (FVMCC_BaseBC.ci:144-187 executeOnTrs -> setGhostState). */
typedef struct cfb200_bc {
  int32_t type;
  int32_t mask[CFB200_MAXEQ];
  double value[CFB200_MAXEQ];
  int32_t zeroGrad[CFB200_MAXEQ];   /* ZeroGradientFlags, FVMCC_ComputeRHS.cxx:186-188 */
} cfb200_bc;

/* DerivativeComputer of the diffusive flux: DiamondVolume2D/3D (DiamondVolume{2,3}DDerivative.cxx, the default) or
 * Corrected2D/3D (CorrectedDerivative2D.cxx:91-430, CorrectedDerivative3D.cxx:54-176) */
enum cfb200_derivative { CFB200_DV_DIAMOND = 0, CFB200_DV_CORRECTED = 1,
                         CFB200_DV_STATEDIFF = 2 /* "StateDiff" plugins/FiniteVolume/StateDiffDerivative.cxx:56-131 */,
                         CFB200_DV_CORRECTED_GG = 3 /* "CorrectedGG2D/3D" plugins/FiniteVolume/CorrectedDerivativeGG2D.cxx,
                                                       CorrectedDerivativeGG3D.cxx: the Corrected face gradient with Green-Gauss
                                                       cell gradients over the nodal states; needs cfb200_mesh::cellNodes */ };

/* Everything ComputeRHS reads from the .CFcase and from SubSystemStatus (SURVEY.md §8b "Scalars crossing per call").
 * The per-call scalars (iter, residual, resFactor, cfl, computeJacobian) may be changed between calls with
 * cfb200_set_step. */
typedef struct cfb200_config {
  int32_t dim;              /* 2 | 3 */
  int32_t nEqs;             /* dim + 2 */
  int32_t updateVar, fluxSplitter, polyRec, limiter, diffusive, viscLaw;
  int32_t computeJacobian;  /* 0: ComputeRHS only; 1: doComputeJacobian() with "NumJacob" (finite differences); 2: "NumJacobAnalytic"
                             * (FVMCC_ComputeRhsJacobAnalytic.cxx: analytic AUSM+up convective blocks on interior faces, Cons) */
  int32_t useFullStencil;   /* Venktn2D.cxx:60 */
  int32_t freezeLimiter;    /* PolyReconstructor.ci:66 */
  int32_t stopLimiting;     /* FVMCC_PolyRec.cxx:52 */
  int32_t isAdimensional;   /* NavierStokesPhysicalModel.ci:137 */
  int32_t freezeDiffCoeff;  /* FVMCC_ComputeRhsJacob option FreezeDiffCoeff */
  int32_t iter;             /* SubSystemStatus::getNbIter() */
  int32_t limitIter;        /* PolyReconstructor.ci:57 */
  double gamma, R, pOffset; /* EulerTerm.cxx:46-50, EulerTerm.hh:199-212 */
  double refValues[CFB200_MAXEQ]; /* PhysicalModel refValues -> NumericalJacobian (RoeFlux.cxx:113-115) */
  double fdTol;             /* NumericalJacobian.cxx:31 */
  double resFactor;         /* getMethodData().getResFactor() */
  double residual;          /* SubSystemStatus::getResidual() */
  double limitRes;          /* PolyReconstructor.ci:60 */
  double gradientFactor;    /* PolyReconstructor.ci:63 */
  double coeffEps, length;  /* Venktn2D.cxx:58,62 */
  double magnitudeValues[CFB200_MAXEQ]; /* Venktn2D.cxx:96-106 */
  double diffRedCoeff;      /* RoeFlux.cxx:62 */
  double coeffTau, coeffQ, cpOverPrandtl, muRef, tempRef, pRef, artDiffCoeff; /* NSTerm / NavierStokesPhysicalModel.ci:69-170 */
  double viscRef, viscTRef, viscSuthConst, fixedVisc;                          /* *SutherlandDynViscosity */
  double cfl;
  /* AUSM family: AUSMFlux.ci:55 choiceA12 (1 Liou critical speed, 2 arithmetic, 3 geometric mean of a_L,a_R);
   * AUSMPlusFlux.ci:17-23 beta, alpha; AUSMPlusUpFlux.ci:20-33 coeffKu, coeffKp, sigma, machInf */
  int32_t choiceA12;
  int32_t hllUseAlpha;      /* HLLFlux.cxx:48 UseAlpha */
  double ausmBeta, ausmAlpha;
  double ausmKu, ausmKp, ausmSigma, machInf;
  /* physical model: CFB200_PM_EULER (Euler / Navier-Stokes, nEqs = dim + 2) or CFB200_PM_MHD3D_PROJECTION
   * (plugins/MHD/MHD3DProjectionCons.cxx, dim 3, nEqs 9: rho, rho u, rho v, rho w, Bx, By, Bz, rho E, phi) */
  int32_t physicalModel;
  int32_t entropyFixID;     /* 0: "Roe" (RoeFlux.cxx); 1 | 2: "RoeEntropyFix" entropyFixID (RoeEntropyFixFlux.cxx:35,57-90) */
  double refSpeed;          /* MHD3DProjection.ConvTerm.refSpeed */
  double coeffSigmoid;      /* AUSMFlux.ci:64 coeffSigmoid (default 1e-8), analytic Jacobian only */
  int32_t derivativeStrategy; /* cfb200_derivative: CellCenterFVM.Data.DerivativeStrategy (CellCenterFVMData.cxx:298) */
  /* 1: cfb200_compute_rhs also applies the pseudo-steady time term (what cfb200_time_rhs does) inside its residual and
   * diagonal gathers: the diagonal blocks and rhs are written once instead of being re-read and re-written by a third
   * kernel.  Same operations in the same order: the results are bit-identical to compute_rhs followed by time_rhs.
   * Not available with timeUseGlobalDT (the global minimum needs the complete updateCoeff first). */
  int32_t fuseTimeRhs;
  /* FVMCC_StdComputeTimeRhs / FVMCC_PseudoSteadyTimeRhs options (FVMCC_StdComputeTimeRhs.cxx:42-57,
   * FVMCC_PseudoSteadyTimeRhs.cxx:40-46,150-175) */
  int32_t timeZeroDiag[CFB200_MAXEQ]; /* zeroDiagValue: equations whose time term is dropped (MHD phi equation) */
  int32_t timeAnalyticalMatrix;       /* useAnalyticalMatrix: dU/dP from Euler2DPuvtToConsInPuvt / Euler3DPvtToConsInPvt */
  int32_t timeUseGlobalDT;            /* useGlobalDT: diag = V / min_cells(V/updateCoeff*CFL) */
  /* 1: cfb200_set_mesh renumbers the cells along a space-filling curve (Morton order of the centres) and sorts the
   * interior faces by their cells, as a pure permutation INSIDE the engine: storage and kernel access order change, every
   * sum keeps the order the reference's numbering defines (FVMCC_MeshDataBuilder.cxx:129-365 IDs stay the parity key) and
   * everything crossing this ABI (states, rhs, updateCoeff, fields, BSR rows / columns, halo lists) stays in the caller's
   * numbering.  For meshes whose numbering has no locality (the gathers then miss L2). */
  int32_t renumber;
  double dt;                          /* SubSystemStatus::getDT(): > 0 (unsteady) -> diag = V/dt, else updateCoeff/CFL */
  /* Venkatakrishnan variants: isMFMHD = the "implementation from the Venkatakrishnan paper" branch (Venktn2D.cxx /
   * Venktn3D.cxx:125-156, normalised by magnitudeValues); CFB200_LIM_VENKTN_STRICT = "Venktn3DStrict"
   * (plugins/FiniteVolumeMHD/Venktn3DStrict.cxx:70-260: that branch with the density-ratio weight strictCoeff on variables
   * 0-3 and 7 and the optional psiMinEqual1 on the others; it has no other branch, so isMFMHD must be set with it) */
  int32_t venktnIsMFMHD;
  int32_t psiMinEqual1;
  double strictCoeff;
  int32_t roeSaOldStencil;            /* RoeSA option "oldStencil" (RoeSAFlux.cxx:44-45,171): every face of the two cells,
                                       * not only those sharing a node with the face */
  /* MHD3DProjection background potential field B0 (MHDTerm option potentialBType, MHDTerm.cxx:25,38-40): 0 = none,
   * 1 = "Dipole" at the origin with moment (mX, mY, mZ) (MHD3DProjectionVarSet.cxx:479-495).  B0 enters the wave speeds of
   * the centred splitters and the update coefficient (computeEigenValues :499-548, getMaxEigenValue :398-436), evaluated at
   * the face quadrature point (FVMCC_PolyRec.cxx:170-185); the flux itself has no B0 term (computeFlux :42-78).
   * "PFSS" needs the reference's own spherical-harmonics data and is not provided. */
  int32_t mhdPotentialBType;
  double mhdDipole[3];
  int32_t noLimiterVar[CFB200_MAXEQ]; /* Venktn2D option "2DNoLimiterID" (Venktn2D.cxx:40-41,253-258): 1 = the limiter of this
                                       * variable is forced to 1 */
  int32_t correctedGGUseWeights;      /* CorrectedGG2D/3D option "useWeights" (CorrectedDerivativeGG2D.cxx:60-61,72-75): inverse
                                       * distance weights in the least squares of the boundary faces (default: none) */
  /* Data.SourceTerm (FVMCC_ComputeRHS::computeSourceTerm, FVMCC_ComputeRHS.cxx:538-600: added to the rhs of a cell, times
   * resFactor, when the face loop first touches the cell; FD Jacobian of the source on the diagonal block,
   * FVMCC_ComputeRhsJacob.cxx:347-365,391-404, .hh:96-105).  CFB200_ST_MHD_CONS_ACA = "MHDConsACAST"
   * (FiniteVolumeMHD/MHDConsACASourceTerm.cxx:164-211) with its solar terms off (gravity = Manchester = RadiativeLossTerm = 0,
   * the defaults: 26 of the reference's 30 cases that name it): the damping term of the divergence-cleaning potential,
   * S[phi] = -(refSpeed^2 / dissipCoeff^2) phi V for MHDTerm correctionType "Mixed" (mhdDissipCoeff = the model's
   * dissipCoeff), zero for "Hyperbolic" (mhdDissipCoeff = 0).  MHD3DProjection only. */
  int32_t sourceTerm;
  double mhdDissipCoeff;
} cfb200_config;

/* Flat mesh data = the contents of the reference's sockets and connectivity tables (SURVEY.md §8b "Data ownership").
 * Faces are in the reference's LOOP order: InnerFaces TRS, then each boundary TRS; faceR of a boundary face is
 * nCells + ghost index.  State-indexed arrays hold the cells first, then the ghost states. */
typedef struct cfb200_mesh {
  int32_t dim, nCells, nGhosts, nNodes, nFaces, nInner;
  const double* nodes;         /* [nNodes*dim]                     socket "nodes" */
  const double* centres;       /* [(nCells+nGhosts)*dim]           State::getCoordinates() of states, gstates */
  const int32_t* faceL;        /* [nFaces]                          face->getState(0)->getLocalID() */
  const int32_t* faceR;        /* [nFaces] */
  const int32_t* faceNodePtr;  /* [nFaces+1] */
  const int32_t* faceNodes;
  const double* normals;       /* [nFaces*dim]  socket "normals" (area weighted, outward from L) */
  const double* areas;         /* [nFaces]      socket "faceAreas" */
  const double* volumes;       /* [nCells]      socket "volumes" */
  const int32_t* cellFacePtr;  /* [nCells+1]    connectivity "cellFaces" */
  const int32_t* cellFaces;
  const int32_t* stencilPtr;   /* [nCells+1]    socket "stencil" */
  const int32_t* stencil;
  const int32_t* nodeNbrPtr;   /* [nNodes+1]    NodalStatesExtrapolator neighbours */
  const int32_t* nodeNbr;
  const double* nodeW;         /*               DistanceBasedExtrapolator weights */
  const double* cvol;          /* [nFaces]      DiamondVolume3DDerivative control volumes (3D) */
  const uint8_t* cvSign;       /* [nFaces]      _changeNormalSign */
  const int32_t* faceBC;       /* [nFaces]      -1 inner, else boundary-condition index */
  const uint8_t* updatable;    /* [nCells]      State::isParUpdatable() */
  const int32_t* cellNodePtr;  /* [nCells+1]    connectivity "cellNodes_InnerCells" (cell-local node order); optional (NULL) */
  const int32_t* cellNodes;    /*               unless derivativeStrategy = CFB200_DV_CORRECTED_GG */
} cfb200_mesh;

/* Device view of the block Jacobian (LSSMatrix replacement; BlockAccumulator.hh:289-296 sub-block layout [ib][jb]). */
typedef struct cfb200_bsr {
  int32_t nRows, nnzb, blockSize;
  const int32_t* rowPtr;   /* device, [nRows+1] */
  const int32_t* colInd;   /* device, [nnzb], ascending inside each row */
  const double* val;       /* device, [nnzb*blockSize*blockSize], row-major blocks */
} cfb200_bsr;

typedef struct cfb200_handle cfb200_handle;

int cfb200_abi_version(void);
/* number of usable CUDA devices (0 without a GPU; never throws) */
int cfb200_device_count(void);

/* diagnostic: compares the kernels' branch-free FP64 divide / square root (csrc/cfb_math.cuh) with the compiler's IEEE
 * `/` and sqrt() on n pseudo-random operand pairs with binary exponents in [-exponentSpan, exponentSpan] (<= 480),
 * zeros and exact quotients included; *mismatches must come back 0 */
int cfb200_selftest_divsqrt(int device, int64_t n, uint64_t seed, int exponentSpan, int64_t* mismatches);

/* replaces: FVMCC_ComputeRHS constructor/configure/setup (plugins/FiniteVolume/FVMCC_ComputeRHS.cxx:46-128,329-440) */
int cfb200_create(const cfb200_config* cfg, int device, cfb200_handle** out);
/* replaces: FVMCC_ComputeRHS::unsetup */
void cfb200_destroy(cfb200_handle* h);
const char* cfb200_last_error(const cfb200_handle* h);

/* replaces: the command's needsSockets()/setup() reads of nodes, normals, faceAreas, volumes, stencil, cellFaces
 * (FVMCC_ComputeRHS.cxx:33-44,329-440; LeastSquareP1Setup.cxx:84-142).  Uploaded once; builds the device SoA,
 * the gather tables (cell->faces in loop order, LSQ edge order, BSR slots) and the LSQ normal matrices. */
int cfb200_set_mesh(cfb200_handle* h, const cfb200_mesh* mesh, int nBC, const cfb200_bc* bcs);
/* per-call scalars: resFactor, iter, residual, cfl, computeJacobian, freezeLimiter ... (whole struct is re-read) */
int cfb200_set_step(cfb200_handle* h, const cfb200_config* cfg);

/* states socket: AoS [nCells][nEqs] as in socket_states.getGlobalArray()->ptr() (FVMCC_ComputeRHSCell.cu:743) */
int cfb200_set_states(cfb200_handle* h, const double* U, int64_t nCells);
int cfb200_get_states(cfb200_handle* h, double* U, int64_t nCells);
/* pastStates socket (NewtonIterator::prepare) for the pseudo-time term */
int cfb200_set_past_states(cfb200_handle* h, const double* U, int64_t nCells);
int cfb200_backup_past_states(cfb200_handle* h);   /* device copy states -> pastStates */
/* Position-dependent Dirichlet data (SuperInlet.cxx:156-209 evaluates its parsed Def function at the boundary point of
 * every face; a time-dependent Def is re-sent before the evaluation): values[iFace*q + v] for the faces of boundary
 * condition record bcIndex in face-loop order replace cfb200_bc::value of a CFB200_BC_DIRICHLET record (the mask stays);
 * NULL reverts the record to its constant.  Call after cfb200_set_mesh. */
int cfb200_set_bc_face_values(cfb200_handle* h, int bcIndex, const double* values);
/* replaces: NodalExtrapolation = DistanceBasedGMove (plugins/FiniteVolume/DistanceBasedExtrapolatorGMove.cxx:92-118 setup,
 * :140-173 applyBC / extrapolate): after the distance-weighted sum, the nodal values idx[i] (option `ValuesIdx`) of the
 * flagged nodes are overwritten by values[i] (option `Values`).  nodeFlag[nNodes] = 1 for the nodes that belong to the TRSs
 * named by option `TRSName` (a node belongs to the first TRS of the priority list that holds it: socket `trsID`).  n <= nEqs.
 * nodeFlag = NULL or n = 0 reverts to the plain "DistanceBased" extrapolation.  Not provided: `RadEquilibrium`, nodal values
 * interpolated from a file (`NodalValuesIDFlags`), `NbIterAdiabatic`.  Call after cfb200_set_mesh. */
int cfb200_set_nodal_prescribed(cfb200_handle* h, const uint8_t* nodeFlag, int32_t n, const int32_t* idx, const double* values);

/* limiter / updateCoeff sockets are stateful across calls (historical limiter; updateCoeff is zeroed by the updater) */
int cfb200_set_limiter(cfb200_handle* h, const double* limiter /* AoS [nCells][nEqs] */);
int cfb200_set_update_coeff(cfb200_handle* h, const double* updateCoeff /* [nCells], NULL = zero */);

/* replaces: CellCenterFVM::applyBC (CellCenterFVM.cxx:464) -> FVMCC_BaseBC::executeOnTrs (FVMCC_BaseBC.ci:144-187) */
int cfb200_apply_bc(cfb200_handle* h);

/* replaces: FVMCC_ComputeRHS::execute (FVMCC_ComputeRHS.cxx:139-325) and, when cfg.computeJacobian,
 * FVMCC_ComputeRhsJacob (FVMCC_ComputeRhsJacob.cxx:126-443).  Runs applyBC first when applyBC != 0.
 * Results stay on the device; rhs/updateCoeff are additionally copied to the host when the pointers are non-NULL
 * (callers UpdateSol.cxx:83 / FVMCC_PseudoSteadyTimeRhs.cxx:106 read them right after). */
int cfb200_compute_rhs(cfb200_handle* h, int applyBC, double* rhs /* [nCells*nEqs] or NULL */,
                       double* updateCoeff /* [nCells] or NULL */);

/* replaces: FVMCC_PseudoSteadyTimeRhs::execute (FVMCC_PseudoSteadyTimeRhs.cxx:106-262), steady branch */
int cfb200_time_rhs(cfb200_handle* h, double* rhs /* or NULL */);

/* replaces: ForwardEuler UpdateSol::execute + ComputeL2Norm (UpdateSol.cxx:83-193, ComputeL2Norm.cxx:81-143):
 * U += CFL/updateCoeff * rhs on the device, updateCoeff = 0, returns log10 L2 norm of dU[monitorVar] */
int cfb200_update_sol_fwd_euler(cfb200_handle* h, double cfl, int monitorVar, double* log10Residual);

/* replaces: NewtonMethod StdUpdateSol::execute + ComputeL2Norm (StdUpdateSol.cxx:110-141,256; NewtonIterator.cxx step):
 * U += alpha[iEq] * dU for the updatable states, dU = 0 for the others, updateCoeff = 0; returns log10 L2 norm of
 * dU[monitorVar].  dU [nCells*q] is the solution of the linear system in the rhs socket layout; NULL = the increment is
 * already in the device rhs buffer (a device-side solver wrote it through cfb200_bsr_view's rhs pointer).
 * alpha [q] = StdUpdateSol.Relaxation, NULL = 1. */
int cfb200_update_sol_newton(cfb200_handle* h, const double* dU, const double* alpha, int monitorVar,
                             double* log10Residual);

/* device-resident results */
int cfb200_bsr_view(cfb200_handle* h, cfb200_bsr* out);
int cfb200_bsr_pattern(cfb200_handle* h, int32_t* rowPtr /* [nCells+1] */, int32_t* colInd /* [nnzb] */);
int64_t cfb200_bsr_nnzb(cfb200_handle* h);
int cfb200_get_bsr_values(cfb200_handle* h, double* val /* [nnzb*q*q] */);
/* scalar CSR view for Paralution (ParalutionMatrix.hh:144 SetDataPtrCSR): rows of a cell one after the other */
int cfb200_export_csr(cfb200_handle* h, int32_t* rowOff, int32_t* col, double* val);
/* the same scalar CSR written by a kernel into caller-provided DEVICE arrays (no download of the values; any array may be
 * NULL): rowOff [nCells*q + 1], col / val [nnzb*q*q]; index64 != 0 -> 64-bit rowOff / col (above 2^31 scalar entries).
 * replaces: the host assembly behind ParalutionMatrix::createSeqBAIJ / SetDataPtrCSR (plugins/Paralution/ParalutionMatrix.cu:48-105) */
int cfb200_export_csr_device(cfb200_handle* h, void* rowOff, void* col, double* val, int index64);
/* replaces: LinearSystemSolver::solveSys with KSPGMRES + block-Jacobi (plugins/Petsc/StdParSolveSys.cxx, PetscLSSData.cxx:60-100:
 * NbKrylovSpaces 30, MaxIter 50, RelativeTolerance 1e-5) for a single partition, entirely on the device: J x = rhs with the
 * right-hand side taken from the device rhs buffer and the solution written back into it (where StdUpdateSol reads dU:
 * cfb200_update_sol_newton(h, NULL, ...)).  Deterministic (ordered sums, no atomics). */
typedef struct cfb200_gmres_params {
  int32_t restart;     /* Krylov vectors per cycle (PETSc NbKrylovSpaces) */
  int32_t maxIter;     /* total iterations */
  int32_t precond;     /* 0 none, 1 block Jacobi (inverted diagonal q x q blocks), right preconditioning */
  int32_t reserved;
  double relTol, absTol;
} cfb200_gmres_params;
int cfb200_solve_gmres(cfb200_handle* h, const cfb200_gmres_params* p, int32_t* iterations, double* relResidual);
/* named fields, reference (AoS, reference IDs) layout: "rhs","updateCoeff","uX","uY","uZ","limiter","nstates","gstates";
 * "divB" [nCells] (MHD3DProjection only): the divergence diagnostic of FVMCC_ComputeRHSMHD.ci:160-203, formed from the
 * face fluxes of the last evaluation */
int cfb200_get_field(cfb200_handle* h, const char* name, double* out);
const double* cfb200_device_ptr(cfb200_handle* h, const char* name);
/* status of the last evaluation: 0 or CFB200_ERR_BAD_VALUE */
int cfb200_status(cfb200_handle* h);

/* per-kernel timing of the last compute call, in ms (CUDA events on the engine's stream; off by default):
 * fills up to cap entries, returns the count */
int cfb200_enable_timing(cfb200_handle* h, int on);
int cfb200_last_timings(cfb200_handle* h, const char** names, float* ms, int cap);
/* stream the engine launches on (cudaStream_t as void*) */
void* cfb200_stream(cfb200_handle* h);

/* ---- multi-GPU (one partition per GPU).  replaces: DataHandle<State*,GLOBAL>::synchronize()
 * (src/Common/MPI/MPICommPattern.hh:1735-1782) for the states array ---------------------------------------- */
int cfb200_nccl_unique_id(void* id128 /* 128 bytes out: ncclUniqueId made on one rank, handed to all by the host code */);
int cfb200_comm_init(cfb200_handle* h, int rank, int nRanks, const void* id128);
/* sendPtr/recvPtr: CSR over ranks [nRanks+1]; sendCells: local ids of the updatable cells to pack for each peer;
 * recvCells: local ids of the overlap cells each peer fills.  Both sides list the cells of a pair in the same
 * (global id) order.  nIndependent: the first nIndependent local cells have no received cell in their reconstruction
 * stencil (their gradients are computed while an exchange is in flight); 0 = no overlap of compute and exchange */
int cfb200_set_halo(cfb200_handle* h, const int32_t* sendPtr, const int32_t* sendCells,
                    const int32_t* recvPtr, const int32_t* recvCells, int32_t nIndependent);
/* the independent cells as a LIST instead of a leading range (overrides nIndependent): the local numbering is then free.
 * With local cells in ascending global ID every gather of an updatable cell runs in the order of the single-domain run, so
 * a partitioned evaluation is bit-identical to it (tests/test_gpu_multi.py, bench.py `parity`). */
int cfb200_set_independent_cells(cfb200_handle* h, const int32_t* cells, int32_t n);
/* starts the exchange (pack kernel, NCCL send/recv group and unpack kernel on a communication stream) and returns;
 * the next cfb200_compute_rhs overlaps it with the work that does not depend on received states */
int cfb200_halo_begin(cfb200_handle* h);
/* begin + wait: the overlap states are in place when it returns */
int cfb200_halo_exchange(cfb200_handle* h);
/* sum over ranks of n <= 64 host doubles (residual norms; replaces the MPI_Allreduce of ComputeL2Norm.cxx:127-137) */
int cfb200_allreduce_sum(cfb200_handle* h, double* v, int n);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The residual norm of
 * cfb200_update_sol_newton / ComputeL2Norm without a handle and WITHOUT ANY ALLOCATION (no cudaMalloc / cudaFree, no
 * allocating library call): d_out[0] = sum over the updatable cells of dU[cell*nEqs + monitorVar]^2, formed with the
 * reduction shape of cfb200_update_sol_newton (same bits); d_out[1] = log10(sqrt(d_out[0])) (-DBL_MAX for a zero sum).
 *   d_dU        device, [nCells*nEqs] increments in the rhs socket layout;   d_updatable  device, [nCells] or NULL = all
 *   d_out       device, 2 doubles, written by the last kernel;   cudaStream  cudaStream_t the kernels are launched on
 *               (NULL = default stream); the call does not synchronise
 *   err/errCap  optional buffer for the error text
 * The block partial sums live in a file-scope device array of the library (131 200 bytes on every GPU that loads it, held
 * until exit), which fixes the limit: nCells <= CFB200_NOALLOC_MAX_CELLS, beyond it the call returns
 * CFB200_ERR_INVALID_ARG with a message.  Every call on a device shares that array and it keeps the last call's
 * contents (each call overwrites every entry it reads, nothing needs clearing): two calls that overlap on the GPU (different
 * streams, no ordering) race; order them.  nCells = 0 is a set-up call: it resolves the array's device address for
 * `device` (done once per device, on the first call) and launches nothing - make it before a graph capture. */
#define CFB200_NOALLOC_MAX_CELLS 16777216
int cfb200_l2_norm_noalloc(int device, const double* d_dU, const uint8_t* d_updatable, int64_t nCells, int32_t nEqs,
                           int32_t monitorVar, void* cudaStream, double* d_out, char* err, int32_t errCap);

/* ---- host-side mesh-data construction (replaces FVMCC_MeshDataBuilder.cxx:91-499, ComputeFaceNormals*P1,
 * FVMCCGeoDataComputer.ci:61-89, ComputeDummyStates.cxx, ComputeFaceVertexNeighborsPlusGhost.cxx,
 * NodalStatesExtrapolator.ci:134-237, DistanceBasedExtrapolator.ci:60-125, DiamondVolume3DDerivative.cxx:413-563,
 * FVMCCSparsity.cxx:47-106) for the stand-alone harness; under COOLFluiD the plugin reads the sockets instead. */
typedef struct cfb200_meshdata cfb200_meshdata;
cfb200_meshdata* cfb200_mesh_build(int dim, int nNodes, const double* nodes, int nCells, const int32_t* cellNodePtr,
                                   const int32_t* cellNodes, const int32_t* cellShape /* 0 tri,1 quad,2 tet,3 hex */,
                                   int nTRS, const int32_t* trsFacePtr, const int32_t* trsFaceCell,
                                   const int32_t* trsFaceNodePtr, const int32_t* trsFaceNodes);
/* why the calling thread's last cfb200_mesh_build returned NULL ("" after a success): unsupported element, a boundary face
 * that matches no cell face, a table beyond the 32-bit offsets of the flat mesh struct: the stencil of more than about 35 M tetrahedra, ... */
const char* cfb200_mesh_build_error(void);
/* sizes: nFaces, nInner, nGhosts, len(faceNodes), len(cellFaces), len(stencil), len(nodeNbr), nnzb */
void cfb200_mesh_sizes(const cfb200_meshdata* m, int64_t* out8);
int cfb200_mesh_get(const cfb200_meshdata* m, const char* name, void* dst);
const void* cfb200_mesh_ptr(const cfb200_meshdata* m, const char* name);
void cfb200_mesh_view(const cfb200_meshdata* m, cfb200_mesh* out);
void cfb200_mesh_set_updatable(cfb200_meshdata* m, const uint8_t* updatable);
/* replaces: the ComputeStencil strategy named by `Setup1.stencil` (LeastSquareP1Setup.cxx; providers "FaceVertex"
 * ComputeFaceVertexNeighbors.cxx:45-129, "FaceEdge" ComputeFaceEdgeNeighbors.cxx:44-150, "Face" ComputeFaceNeighbors.cxx:46-108).
 * cfb200_mesh_build leaves "FaceVertexPlusGhost" (ComputeFaceVertexNeighborsPlusGhost.cxx:46-230); this call replaces
 * stencilPtr / stencil of *m by one of the others (FaceVertex / FaceEdge in ascending cell ID, Face in cellFaces order; FaceEdge
 * and Face list each edge once, at its lower cell, as the reference does).  Returns CFB200_ERR_INVALID_ARG for any other kind.
 * Under COOLFluiD the plugin reads whatever stencil the configured strategy left in socket "stencil". */
enum { CFB200_STENCIL_FACE_VERTEX_PLUS_GHOST = 0, CFB200_STENCIL_FACE_VERTEX = 1, CFB200_STENCIL_FACE_EDGE = 2, CFB200_STENCIL_FACE = 3 };
int cfb200_mesh_set_stencil(cfb200_meshdata* m, int kind);
void cfb200_mesh_free(cfb200_meshdata* m);
/* cvol / cvSign of a mesh assembled by the caller (DiamondVolume3DDerivative.cxx:413-563; zeros in 2D); the other
 * fields of *m must be set */
void cfb200_mesh_diamond_volumes(const cfb200_mesh* m, double* cvol /* [nFaces] */, uint8_t* cvSign /* [nFaces] */);

#ifdef __cplusplus
}
#endif
#endif
