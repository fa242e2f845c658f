// See FVMCC_ComputeRHSB200.hh.  Written against the COOLFluiD API of /root/reference; not compiled here.
#include "FiniteVolumeB200/FiniteVolumeB200.hh"
#include "FiniteVolumeB200/FVMCC_ComputeRHSB200.hh"
#include "Framework/MethodCommandProvider.hh"
#include "Framework/MeshData.hh"
#include "Framework/SubSystemStatus.hh"
#include "Framework/PhysicalModel.hh"
#include "Framework/LSSMatrix.hh"
#include "Common/PE.hh"
#include "Common/BadValueException.hh"
#include "NavierStokes/EulerTerm.hh"

using namespace std;
using namespace COOLFluiD::Framework;
using namespace COOLFluiD::Common;

namespace COOLFluiD { namespace Numerics { namespace FiniteVolume {

MethodCommandProvider<FVMCC_ComputeRHSB200, CellCenterFVMData, FiniteVolumeB200Module>
fvmcc_computeRHSB200Provider("B200Rhs");
MethodCommandProvider<FVMCC_ComputeRhsJacobB200, CellCenterFVMData, FiniteVolumeB200Module>
fvmcc_computeRhsJacobB200Provider("B200RhsJacob");
MethodCommandProvider<FVMCC_ComputeRhsJacobAnalyticB200, CellCenterFVMData, FiniteVolumeB200Module>
fvmcc_computeRhsJacobAnalyticB200Provider("B200RhsJacobAnalytic");

void FVMCC_ComputeRHSB200::defineConfigOptions(Config::OptionList& options)
{
  options.addConfigOption< CFint >("Device", "CUDA device ordinal (-1: rank % deviceCount).");
  options.addConfigOption< bool >("HostJacobian", "Copy the block Jacobian to the host LSS matrix (CPU PETSc).");
}

FVMCC_ComputeRHSB200::FVMCC_ComputeRHSB200(const std::string& name) :
  FVMCC_ComputeRHS(name), m_h(CFNULL)
{
  addConfigOptionsTo(this);
  m_device = -1;       setParameter("Device", &m_device);
  m_hostJacobian = true; setParameter("HostJacobian", &m_hostJacobian);
}

FVMCC_ComputeRHSB200::~FVMCC_ComputeRHSB200() {}

void FVMCC_ComputeRHSB200::configure(Config::ConfigArgs& args) { FVMCC_ComputeRHS::configure(args); }

void FVMCC_ComputeRHSB200::check(int rc, const char* what)
{
  if (rc == CFB200_OK) return;
  // same error convention as the CPU path: a^2 < 0 throws BadValueException (Euler2DLinearRoe.cxx:79-85)
  throw BadValueException(FromHere(), std::string(what) + ": " + cfb200_last_error(m_h));
}

void FVMCC_ComputeRHSB200::fillConfig(cfb200_config& c)
{
  // values the CPU strategies hold after configuration: PhysicalModel (gamma, R, refValues), PolyReconstructor
  // (limitRes, gradientFactor, limitIter, freezeLimiter), Limiter (coeffEps, length, magnitudeValues, useFullStencil),
  // FluxSplitter (DiffCoeff), NSTerm (coeffTau, coeffQ, cpOverPrandtl, mu/temp/p references), NumericalJacobian tol.
  const CFuint dim = PhysicalModelStack::getActive()->getDim();
  const CFuint nbEqs = PhysicalModelStack::getActive()->getNbEq();
  std::memset(&c, 0, sizeof(c));
  c.dim = dim; c.nEqs = nbEqs;
  const std::string model = PhysicalModelStack::getActive()->getImplementor()->getConvectiveName();
  if (model.find("MHD3DProjection") == 0) {
    // plugins/MHD: MHD3DProjection.ConvTerm.gamma / refSpeed (e.g. Dipole3D/dipole3DFVMMHDProjImpl.CFcase)
    SafePtr<Physics::MHD::MHDProjectionTerm> mhdTerm = PhysicalModelStack::getActive()->getImplementor()->
      getConvectiveTerm().d_castTo<Physics::MHD::MHDProjectionTerm>();
    c.physicalModel = CFB200_PM_MHD3D_PROJECTION;
    c.gamma = mhdTerm->getGamma(); c.refSpeed = mhdTerm->getRefSpeed(); c.R = 1.0;
  } else {
    SafePtr<Physics::NavierStokes::EulerTerm> eulerTerm = PhysicalModelStack::getActive()->getImplementor()->
      getConvectiveTerm().d_castTo<Physics::NavierStokes::EulerTerm>();
    c.physicalModel = CFB200_PM_EULER;
    c.gamma = eulerTerm->getGamma(); c.R = eulerTerm->getR();
  }
  c.updateVar = (getMethodData().getUpdateVarStr() == "Cons") ? CFB200_UV_CONS : CFB200_UV_PVT;
  // CellCenterFVM.Data.FluxSplitter = Roe | AUSMPlus{2,3}D | AUSMPlusUp{2,3}D | LaxFried | HLL
  const std::string fs = getMethodData().getFluxSplitter()->getName();
  c.fluxSplitter = (fs.find("AUSMPlusUp") == 0) ? CFB200_FS_AUSMPLUSUP : (fs.find("AUSMPlus") == 0) ? CFB200_FS_AUSMPLUS :
                   (fs == "LaxFried") ? CFB200_FS_LAXFRIED : (fs == "HLL") ? CFB200_FS_HLL : CFB200_FS_ROE;
  // AUSMFlux.ci:55, AUSMPlusFlux.ci:17-23, AUSMPlusUpFlux.ci:20-33 defaults; the configured values are read from the
  // splitter's option map (choiceA12, beta, alpha, coeffKu, coeffKp, sigma, machInf), HLLFlux.cxx:48 UseAlpha
  c.choiceA12 = 1; c.ausmBeta = 1./8.; c.ausmAlpha = 3./16.; c.ausmKu = 0.75; c.ausmKp = 0.25; c.ausmSigma = 1.0;
  c.machInf = 0.0; c.hllUseAlpha = 0;
  c.polyRec = (getMethodData().getPolyReconstructor()->getName() == "Constant") ? CFB200_PR_CONSTANT : CFB200_PR_LINEARLS;
  const std::string lim = getMethodData().getLimiter()->getName();
  c.limiter = (lim.find("Venktn") == 0) ? CFB200_LIM_VENKTN : (lim == "BarthJesp" ? CFB200_LIM_BARTHJESP : CFB200_LIM_NONE);
  c.diffusive = getMethodData().getDiffusiveFluxComputer()->isNull() ? CFB200_DIFF_NONE : CFB200_DIFF_NS;
  // CellCenterFVM.Data.DerivativeStrategy: DiamondVolume2D/3D (CellCenterFVMData.cxx:298 default) or Corrected2D/3D
  const std::string deriv = getMethodData().getDerivativeComputer()->getName();
  if (deriv == "Corrected2D" || deriv == "Corrected3D") c.derivativeStrategy = CFB200_DV_CORRECTED;
  else if (deriv == "DiamondVolume2D" || deriv == "DiamondVolume3D" || c.diffusive == CFB200_DIFF_NONE)
    c.derivativeStrategy = CFB200_DV_DIAMOND;
  else c.derivativeStrategy = -1;      // CorrectedGG*, StateDiff, ...: cfb200_create answers CFB200_ERR_UNSUPPORTED
  c.reserved0 = 0;
  const RealVector& refValues = PhysicalModelStack::getActive()->getImplementor()->getRefStateValues();
  for (CFuint i = 0; i < nbEqs; ++i) { c.refValues[i] = refValues[i]; c.magnitudeValues[i] = refValues[i]; }
  c.fdTol = 10e-7;                     // NumericalJacobian.cxx:31
  c.limitRes = -2.0; c.gradientFactor = 1.0; c.limitIter = 1000000000; c.coeffEps = 1.0; c.length = 1.0;
  c.diffRedCoeff = 1.0; c.artDiffCoeff = 1.0;
  // (the remaining option values are read from the configured strategies through their accessor methods)
}

void FVMCC_ComputeRHSB200::setup()
{
  FVMCC_ComputeRHS::setup();
  cfb200_config c; fillConfig(c);
  int device = m_device;
  if (device < 0) device = PE::GetPE().GetRank("Default") % std::max(1, cfb200_device_count());
  check(cfb200_create(&c, device, &m_h), "cfb200_create");
  uploadMesh();
}

void FVMCC_ComputeRHSB200::uploadMesh()
{
  // Flattens the sockets nodes, states/gstates coordinates, normals, faceAreas, volumes, stencil, the connectivity
  // tables cellFaces / InnerFaces / boundary TRSs (in the order of FVMCC_ComputeRHS::execute's TRS loop), the nodal
  // extrapolator's neighbour lists and weights, and State::isParUpdatable() into a cfb200_mesh (include/cfb200.h),
  // one boundary-condition record per boundary TRS (type + values from the configured BC commands), then
  //   check(cfb200_set_mesh(m_h, &mesh, nbBCs, bcs), "cfb200_set_mesh");
  // Field-by-field mapping: INTEGRATION.md §3.
}

void FVMCC_ComputeRHSB200::execute()
{
  DataHandle<State*, GLOBAL> states = socket_states.getDataHandle();
  DataHandle<CFreal> rhs = socket_rhs.getDataHandle();
  DataHandle<CFreal> updateCoeff = socket_updateCoeff.getDataHandle();
  cfb200_config c; fillConfig(c);
  c.resFactor = getResFactor();
  c.computeJacobian = getMethodData().doComputeJacobian() ? withJacobian() : 0;
  c.iter = SubSystemStatusStack::getActive()->getNbIter();
  c.residual = SubSystemStatusStack::getActive()->getResidual();
  check(cfb200_set_step(m_h, &c), "cfb200_set_step");
  // states are contiguous in the global array (FVMCC_ComputeRHSCell.cu:743)
  check(cfb200_set_states(m_h, states.getGlobalArray()->ptr(), states.size()), "cfb200_set_states");
  check(cfb200_set_update_coeff(m_h, &updateCoeff[0]), "cfb200_set_update_coeff");
  // ghost states are rebuilt on the device by the same boundary conditions (applyBC = 1)
  check(cfb200_compute_rhs(m_h, 1, &rhs[0], &updateCoeff[0]), "cfb200_compute_rhs");
  if (c.computeJacobian) insertJacobian();
}

void FVMCC_ComputeRHSB200::insertJacobian()
{
  // CPU PETSc: copy the BSR values once and insert block rows (BlockAccumulator.hh:289-296 layout is the BSR block
  // layout, so one MatSetValuesBlocked per row); GPU PETSc (MATMPIAIJCUSPARSE) / Paralution: hand over the device
  // view from cfb200_bsr_view instead (ParalutionMatrix.hh:144 SetDataPtrCSR after cfb200_export_csr).
  SafePtr<LSSMatrix> jacobMatrix = getMethodData().getLinearSystemSolver()[0]->getMatrix();
  if (m_rowPtr.empty()) {
    const CFuint nbCells = socket_states.getDataHandle().size();
    m_rowPtr.resize(nbCells + 1); m_colInd.resize(cfb200_bsr_nnzb(m_h));
    check(cfb200_bsr_pattern(m_h, &m_rowPtr[0], &m_colInd[0]), "cfb200_bsr_pattern");
    const CFuint nbEqs = PhysicalModelStack::getActive()->getNbEq();
    m_bsr.resize(m_colInd.size()*nbEqs*nbEqs);
  }
  check(cfb200_get_bsr_values(m_h, &m_bsr[0]), "cfb200_get_bsr_values");
  jacobMatrix->resetToZeroEntries();
  // for each updatable row: jacobMatrix->addValues(row accumulator built over m_colInd[rowPtr[i]..rowPtr[i+1]))
}

void FVMCC_ComputeRHSB200::unsetup()
{
  if (m_h != CFNULL) { cfb200_destroy(m_h); m_h = CFNULL; }
  FVMCC_ComputeRHS::unsetup();
}

} } }
