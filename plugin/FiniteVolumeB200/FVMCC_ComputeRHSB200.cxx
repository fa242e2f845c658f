// See FVMCC_ComputeRHSB200.hh.  Written against the COOLFluiD API of /root/reference; not compiled here.
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <map>
#include <algorithm>
#include <typeinfo>
#include "FiniteVolumeB200/FiniteVolumeB200.hh"
#include "FiniteVolumeB200/FVMCC_ComputeRHSB200.hh"
#include "Framework/MethodCommandProvider.hh"
#include "Framework/MeshData.hh"
#include "Framework/SubSystemStatus.hh"
#include "Framework/PhysicalModel.hh"
#include "Framework/LSSMatrix.hh"
#include "Common/PE.hh"
#include "Common/BadValueException.hh"
#include "Common/NotImplementedException.hh"
#include "Common/NoSuchStorageException.hh"
#include "NavierStokes/EulerTerm.hh"
#include "Framework/NodalStatesExtrapolator.hh"
#include "Framework/ConnectivityTable.hh"
#include "MathTools/MathFunctions.hh"
#include "Common/DemangledTypeID.hh"
#include "FiniteVolume/FVMCC_BC.hh"
#include "FiniteVolume/FVMCC_PolyRec.hh"
#include "Framework/Limiter.hh"
#include "Framework/FluxSplitter.hh"
#include "Framework/ComputeSourceTerm.hh"
#include "Framework/CFL.hh"
#include "Framework/BlockAccumulator.hh"
#include "Framework/LinearSystemSolver.hh"
#include "Config/ConfigOptionException.hh"
#include "NavierStokes/NSTerm.hh"
#include "MHD/MHDProjectionTerm.hh"
#include <sstream>

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

std::vector<SafePtr<BaseDataSocketSink> > FVMCC_ComputeRHSB200::needsSockets()
{
  std::vector<SafePtr<BaseDataSocketSink> > result = FVMCC_ComputeRHS::needsSockets();
  result.push_back(&socket_stencil);
  result.push_back(&socket_pastStates);      // not essential: only the fused pseudo-steady term reads it
  result.push_back(&socket_trsID);           // not essential: only NodalExtrapolation = DistanceBasedGMove reads it
  return result;
}

void FVMCC_ComputeRHSB200::defineConfigOptions(Config::OptionList& options)
{
  options.addConfigOption< CFint >("Device", "CUDA device ordinal (-1: rank % deviceCount).");
  options.addConfigOption< bool >("HostJacobian", "Copy the block Jacobian to the host LSS matrix (CPU PETSc).");
  options.addConfigOption< CFreal >("ViscRef", "Reference viscosity of the configured viscosity law (NavierStokes.DiffTerm.*.ViscRef).");
  options.addConfigOption< CFreal >("TRef", "Reference temperature of the configured viscosity law.");
  options.addConfigOption< CFreal >("SuthConst", "Sutherland constant of the configured viscosity law.");
  options.addConfigOption< CFreal >("FixedVisc", "Viscosity of a constant viscosity law.");
  options.addConfigOption< bool >("RefreshInletValues", "Re-evaluate the Dirichlet-type boundary commands (SuperInlet ...) on every face before each evaluation: for Def functions of time or iteration.");
  options.addConfigOption< bool >("FuseTimeRhs", "Apply the pseudo-steady time term inside this command (ComputeTimeRHS must then be Null).");
  options.addConfigOption< bool >("useAnalyticalMatrix", "As PseudoSteadyTimeRhs.useAnalyticalMatrix (fused time term).");
  options.addConfigOption< bool >("useGlobalDT", "As PseudoSteadyTimeRhs.useGlobalDT.");
  options.addConfigOption< std::vector<bool> >("zeroDiagValue", "As PseudoSteadyTimeRhs.zeroDiagValue (fused time term).");
}

FVMCC_ComputeRHSB200::FVMCC_ComputeRHSB200(const std::string& name) :
  FVMCC_ComputeRHS(name), socket_stencil("stencil"), socket_pastStates("pastStates", false), socket_trsID("trsID", false),
  m_h(CFNULL)
{
  addConfigOptionsTo(this);
  m_device = -1;       setParameter("Device", &m_device);
  m_hostJacobian = true; setParameter("HostJacobian", &m_hostJacobian);
  m_viscRef = 0.;        setParameter("ViscRef", &m_viscRef);
  m_viscTRef = 0.;       setParameter("TRef", &m_viscTRef);
  m_viscSuthConst = 0.;  setParameter("SuthConst", &m_viscSuthConst);
  m_fixedVisc = 1.;      setParameter("FixedVisc", &m_fixedVisc);
  m_refreshInletValues = false; setParameter("RefreshInletValues", &m_refreshInletValues);
  m_fuseTimeRhs = false; setParameter("FuseTimeRhs", &m_fuseTimeRhs);
  m_timeAnalyticalMatrix = false; setParameter("useAnalyticalMatrix", &m_timeAnalyticalMatrix);
  m_timeUseGlobalDT = false;      setParameter("useGlobalDT", &m_timeUseGlobalDT);
  m_timeZeroDiag = std::vector<bool>(); setParameter("zeroDiagValue", &m_timeZeroDiag);
  std::memset(&m_cfg, 0, sizeof(m_cfg));
}

FVMCC_ComputeRHSB200::~FVMCC_ComputeRHSB200() {}

void FVMCC_ComputeRHSB200::configure(Config::ConfigArgs& args) { FVMCC_ComputeRHS::configure(args); }

void FVMCC_ComputeRHSB200::check(int rc, const char* what)
{
  if (rc == CFB200_OK) return;
  // same error convention as the CPU path: a^2 < 0 throws BadValueException (Euler2DLinearRoe.cxx:79-85)
  throw BadValueException(FromHere(), std::string(what) + ": " + cfb200_last_error(m_h));
}

// ---- option read-back -------------------------------------------------------------------------------------------------
// Every strategy is a Config::ConfigObject whose options are linked to its members (ConfigObject.hh:143-175 setParameter),
// so Option::getValueAsString() (OptionT.hh:52,153) returns the CURRENT value, configured or default, also of members that
// have no accessor (Venktn2D::_coeffEps, AUSMPlusUpFlux::m_coeffKu, ...).  Vectors come back blank separated, bool as 0/1.
namespace {
template <class OBJ> bool hasOpt(OBJ& o, const std::string& name) {
  try { o.getOption(name); return true; } catch (Config::ConfigOptionException&) { return false; }
}
template <class OBJ> std::string optStr(OBJ& o, const std::string& name, const std::string& dflt) {
  try { return o.getOption(name)->getValueAsString(); } catch (Config::ConfigOptionException&) { return dflt; }
}
template <class OBJ> CFreal optReal(OBJ& o, const std::string& name, CFreal dflt) {
  const std::string v = optStr(o, name, "");
  return v.empty() ? dflt : std::strtod(v.c_str(), CFNULL);
}
template <class OBJ> CFint optInt(OBJ& o, const std::string& name, CFint dflt) {
  const std::string v = optStr(o, name, "");
  if (v == "true") return 1;
  if (v == "false") return 0;
  return v.empty() ? dflt : (CFint)std::strtol(v.c_str(), CFNULL, 10);
}
template <class OBJ> std::vector<CFreal> optVec(OBJ& o, const std::string& name) {
  std::vector<CFreal> r;
  std::istringstream is(optStr(o, name, ""));
  std::string tok;
  while (is >> tok) r.push_back((tok == "true") ? 1. : (tok == "false") ? 0. : std::strtod(tok.c_str(), CFNULL));
  return r;
}
}

// The static part of cfb200_config: everything the CPU strategies hold after configuration.  Built ONCE in setup() (the
// strategies are configured by then) and cached in m_cfg; execute() only refreshes the per-call scalars.  Field by field
// this follows include/cfb200.h; INTEGRATION.md section 4 lists option -> accessor -> field.
void FVMCC_ComputeRHSB200::fillConfig(cfb200_config& c)
{
  CellCenterFVMData& md = getMethodData();
  const CFuint dim = PhysicalModelStack::getActive()->getDim();
  const CFuint nbEqs = PhysicalModelStack::getActive()->getNbEq();
  SafePtr<PhysicalModelImpl> pm = PhysicalModelStack::getActive()->getImplementor();
  std::memset(&c, 0, sizeof(c));
  c.dim = dim; c.nEqs = nbEqs;

  // ---- physical model: convective term
  const std::string model = pm->getConvectiveName();
  if (model.find("MHD3DProjection") == 0) {
    // plugins/MHD: MHD3DProjection.ConvTerm.gamma / refSpeed (e.g. Dipole3D/dipole3DFVMMHDProjImpl.CFcase)
    SafePtr<Physics::MHD::MHDProjectionTerm> mhdTerm = pm->getConvectiveTerm().d_castTo<Physics::MHD::MHDProjectionTerm>();
    c.physicalModel = CFB200_PM_MHD3D_PROJECTION;
    c.gamma = mhdTerm->getGamma(); c.refSpeed = mhdTerm->getRefSpeed(); c.R = 1.0;
    // background potential field (MHDTerm.hh:203-232): "Dipole" is evaluated by the engine, anything else but "" is refused
    const std::string pb = mhdTerm->getPotentialBType();
    c.mhdPotentialBType = (pb == "Dipole") ? 1 : (pb.empty() ? 0 : -1);
    c.mhdDipole[0] = mhdTerm->getMX(); c.mhdDipole[1] = mhdTerm->getMY(); c.mhdDipole[2] = mhdTerm->getMZ();
    // Data.SourceTerm = MHDConsACAST (MHDConsACASourceTerm.cxx:164-211): the phi damping of correctionType "Mixed"
    // (MHDTerm.hh:197,238); its solar terms (gravity, Manchester, RadiativeLossTerm, :211) are not provided
    const std::vector<std::string> stNames = md.getSourceTermComputerNames();
    if (stNames.size() > 1 || (stNames.size() == 1 && stNames[0] != "MHDConsACAST" && stNames[0] != "Null"))
      throw Common::NotImplementedException(FromHere(), "B200: Data.SourceTerm other than MHDConsACAST");
    if (stNames.size() == 1 && stNames[0] == "MHDConsACAST") {
      SafePtr<ComputeSourceTerm<CellCenterFVMData> > st = (*md.getSourceTermComputer())[0];
      if (optInt(*st, "gravity", 0) != 0 || optInt(*st, "Manchester", 0) != 0 || optInt(*st, "RadiativeLossTerm", 0) != 0)
        throw Common::NotImplementedException(FromHere(), "B200: MHDConsACAST with gravity / Manchester / RadiativeLossTerm");
      c.sourceTerm = CFB200_ST_MHD_CONS_ACA;
      c.mhdDissipCoeff = (mhdTerm->getCorrectionType() == "Mixed") ? mhdTerm->getDissipationCoefficient() : 0.;
    }
  } else {
    SafePtr<Physics::NavierStokes::EulerTerm> eulerTerm = pm->getConvectiveTerm().d_castTo<Physics::NavierStokes::EulerTerm>();
    c.physicalModel = CFB200_PM_EULER;
    c.gamma = eulerTerm->getGamma();                       // EulerTerm.hh:139
    c.R = eulerTerm->getR();                               // EulerTerm.hh:129 (_RRef)
    c.pOffset = eulerTerm->getPressureFromState(0.);       // EulerTerm.hh:199-202: _p0Inf/refP
    c.tempRef = eulerTerm->getTempRef();                   // EulerTerm.hh:174
    c.pRef = eulerTerm->getPressRef();                     // EulerTerm.hh:179
  }
  c.isAdimensional = pm->isAdimensional() ? 1 : 0;         // PhysicalModelImpl.hh:120
  c.updateVar = (md.getUpdateVarStr() == "Cons") ? CFB200_UV_CONS : CFB200_UV_PVT;
  const RealVector& refValues = pm->getRefStateValues();   // PhysicalModelImpl.hh:108 -> NumericalJacobian (RoeFlux.cxx:113-115)
  for (CFuint i = 0; i < nbEqs; ++i) { c.refValues[i] = refValues[i]; c.magnitudeValues[i] = refValues[i]; }
  c.fdTol = 10e-7;                                         // NumericalJacobian.cxx:31 (no option, no accessor)

  // ---- flux splitter: CellCenterFVM.Data.FluxSplitter = Roe | RoeEntropyFix | RoeSA | AUSMPlus{2,3}D | AUSMPlusUp{2,3}D | LaxFried | HLL
  FluxSplitter<CellCenterFVMData>& fsp = *md.getFluxSplitter();
  const std::string fs = fsp.getName();
  c.fluxSplitter = (fs.find("AUSMPlusUp") == 0) ? CFB200_FS_AUSMPLUSUP : (fs.find("AUSMPlus") == 0) ? CFB200_FS_AUSMPLUS :
                   (fs == "LaxFried") ? CFB200_FS_LAXFRIED : (fs == "HLL") ? CFB200_FS_HLL : (fs == "Centred") ? CFB200_FS_CENTRED :
                   (fs == "Roe" || fs == "RoeEntropyFix") ? CFB200_FS_ROE : (fs == "RoeSA") ? CFB200_FS_ROESA :
                   -1;                                                               // -1: cfb200_create -> UNSUPPORTED
  c.diffRedCoeff = optReal(fsp, "DiffCoeff", 1.0);         // RoeFlux.cxx:34,62 / LaxFriedFlux.cxx:32
  c.entropyFixID = (fs == "RoeEntropyFix") ? optInt(fsp, "entropyFixID", 1) :        // RoeEntropyFixFlux.cxx:35,49
                   (fs == "RoeSA") ? optInt(fsp, "entropyFixID", 0) : 0;             // RoeSAFlux.cxx:41-42
  c.roeSaOldStencil = (fs == "RoeSA") ? optInt(fsp, "oldStencil", 0) : 0;            // RoeSAFlux.cxx:44-45
  c.choiceA12 = optInt(fsp, "choiceA12", 1);               // AUSMFlux.ci:55,73
  c.coeffSigmoid = optReal(fsp, "coeffSigmoid", 1e-8);     // AUSMFlux.ci:64,84
  c.ausmBeta = optReal(fsp, "beta", 1./8.);                // AUSMPlusFlux.ci:38 / AUSMPlusUpFlux.ci:58
  c.ausmAlpha = optReal(fsp, "alpha", 3./16.);             // AUSMPlusFlux.ci:41
  c.ausmKu = optReal(fsp, "coeffKu", 0.75);                // AUSMPlusUpFlux.ci:50
  c.ausmKp = optReal(fsp, "coeffKp", 0.25);                // AUSMPlusUpFlux.ci:52
  c.ausmSigma = optReal(fsp, "sigma", 1.0);                // AUSMPlusUpFlux.ci:54
  c.machInf = optReal(fsp, "machInf", 0.0);                // AUSMPlusUpFlux.ci:56
  c.hllUseAlpha = optInt(fsp, "UseAlpha", 0);              // HLLFlux.cxx:32,48

  // ---- reconstruction + limiter
  FVMCC_PolyRec& rec = *md.getPolyReconstructor();
  c.polyRec = (rec.getName() == "Constant") ? CFB200_PR_CONSTANT : CFB200_PR_LINEARLS;
  c.limitIter = optInt(rec, "limitIter", 1000000000);      // PolyReconstructor.ci:29,57
  c.limitRes = optReal(rec, "limitRes", -2.0);             // PolyReconstructor.ci:31,60
  c.gradientFactor = optReal(rec, "gradientFactor", 1.0);  // PolyReconstructor.ci:33,63 (constant; a Def of the iteration is
                                                           //  re-evaluated per call by refreshStep())
  c.freezeLimiter = optInt(rec, "freezeLimiter", 0);       // PolyReconstructor.ci:35,66
  c.stopLimiting = optInt(rec, "StopLimiting", 0);         // FVMCC_PolyRec.cxx:27,52
  Limiter<CellCenterFVMData>& limr = *md.getLimiter();
  const std::string lim = limr.getName();
  c.limiter = (lim == "Venktn2D" || lim == "Venktn3D") ? CFB200_LIM_VENKTN : (lim.find("BarthJesp") == 0) ? CFB200_LIM_BARTHJESP :    // BarthJesp | BarthJesp2D | BarthJesp3D: one class (BarthJesp.cxx:26-34)
             
              (lim == "Venktn3DStrict") ? CFB200_LIM_VENKTN_STRICT :
              (lim == "Null" || c.polyRec == CFB200_PR_CONSTANT) ? CFB200_LIM_NONE : -1;
  c.coeffEps = optReal(limr, "coeffEps", 1.0);             // Venktn2D.cxx:34,58
  c.length = optReal(limr, "length", 1.0);                 // Venktn2D.cxx:36,62
  c.useFullStencil = optInt(limr, "useFullStencil", 0);    // Limiter.ci:22
  {
    const std::vector<CFreal> mv = optVec(limr, "magnitudeValues");              // Venktn2D.cxx:32,96-106: default = refValues
    if (mv.size() == nbEqs) for (CFuint i = 0; i < nbEqs; ++i) c.magnitudeValues[i] = mv[i];
  }
  {
    const std::vector<CFreal> nl = (lim == "Venktn2D") ? optVec(limr, "2DNoLimiterID") : std::vector<CFreal>();   // Venktn2D.cxx:40-41
    for (CFuint i = 0; i < CFB200_MAXEQ; ++i) c.noLimiterVar[i] = 0;
    for (CFuint i = 0; i < nl.size(); ++i) if (nl[i] >= 0 && nl[i] < nbEqs) c.noLimiterVar[(CFuint)nl[i]] = 1;
  }
  c.venktnIsMFMHD = (c.limiter == CFB200_LIM_VENKTN_STRICT) ? 1 : optInt(limr, "isMFMHD", 0);   // Venktn3D.cxx:125-156
  c.strictCoeff = optReal(limr, "strictCoeff", 1.0);       // Venktn3DStrict.cxx:31,44
  c.psiMinEqual1 = optInt(limr, "psiMinEqual1", 0);        // Venktn3DStrict.cxx:33,47

  // ---- diffusive flux
  c.diffusive = md.getDiffusiveFluxComputer()->isNull() ? CFB200_DIFF_NONE : CFB200_DIFF_NS;
  // CellCenterFVM.Data.DerivativeStrategy: DiamondVolume2D/3D (CellCenterFVMData.cxx:298 default), Corrected2D/3D, CorrectedGG2D/3D or StateDiff
  const std::string deriv = md.getDerivativeComputer()->getName();
  if (deriv == "Corrected2D" || deriv == "Corrected3D") c.derivativeStrategy = CFB200_DV_CORRECTED;
  else if (deriv == "StateDiff") c.derivativeStrategy = CFB200_DV_STATEDIFF;
  else if (deriv == "CorrectedGG2D" || deriv == "CorrectedGG3D") {
    c.derivativeStrategy = CFB200_DV_CORRECTED_GG;
    c.correctedGGUseWeights = optInt(*md.getDerivativeComputer(), "useWeights", 0);    // CorrectedDerivativeGG2D.cxx:60-61,72-75
  }
  else if (deriv == "DiamondVolume2D" || deriv == "DiamondVolume3D" || c.diffusive == CFB200_DIFF_NONE)
    c.derivativeStrategy = CFB200_DV_DIAMOND;
  else c.derivativeStrategy = -1;      // anything else: cfb200_create answers CFB200_ERR_UNSUPPORTED
  c.coeffTau = 1.0; c.coeffQ = 1.0; c.cpOverPrandtl = 1.0; c.muRef = 1.0; c.artDiffCoeff = 1.0;
  if (c.diffusive == CFB200_DIFF_NS) {
    SafePtr<Physics::NavierStokes::NSTerm> ns = pm->getDiffusiveTerm().d_castTo<Physics::NavierStokes::NSTerm>();
    c.coeffTau = ns->getCoeffTau();                        // NSTerm.hh:190
    c.coeffQ = ns->getCoeffQ();                            // NSTerm.hh:199
    c.cpOverPrandtl = ns->getCpOverPrandtl();              // NSTerm.hh:208
    c.artDiffCoeff = ns->getArtDiffCoeff();                // NSTerm.hh:234
    c.muRef = ns->getReferencePhysicalData()[Physics::NavierStokes::NSTerm::MU];   // NavierStokesPhysicalModel.ci:69-170
    // The viscosity law object is private to NSTerm (NSTerm.hh:251, no accessor): its name is NSTerm's "ViscosityLaw" option,
    // its constants are this command's options ViscRef / TRef / SuthConst / FixedVisc (defaults = the reference's air
    // values); setup() then VERIFIES the translated law against NSTerm::getDynViscosityDim and refuses to run otherwise.
    const std::string law = optStr(*ns, "ViscosityLaw", "SimplerSutherland");     // NSTerm.cxx:47-48
    c.viscLaw = (law == "SimplerSutherland") ? CFB200_VISC_SIMPLER_SUTHERLAND : (law == "Sutherland") ? CFB200_VISC_SUTHERLAND :
                (law.find("Fixed") == 0 || law.find("Constant") == 0) ? CFB200_VISC_FIXED : -1;
    c.viscRef = (m_viscRef > 0.) ? m_viscRef : (c.viscLaw == CFB200_VISC_SUTHERLAND ? 18.27e-6 : 0.000001458);
    c.viscTRef = (m_viscTRef > 0.) ? m_viscTRef : (c.viscLaw == CFB200_VISC_SUTHERLAND ? 291.15 : 110.4);
    c.viscSuthConst = (m_viscSuthConst > 0.) ? m_viscSuthConst : 120.0;
    c.fixedVisc = m_fixedVisc;
    if (c.viscLaw >= 0) {
      const CFreal Ts[3] = {0.7*c.tempRef, c.tempRef, 2.5*c.tempRef};
      for (int k = 0; k < 3; ++k) {
        const CFreal T = Ts[k] > 0. ? Ts[k] : 100.*(k + 1);
        const CFreal ref = ns->getDynViscosityDim(c.pRef, T);                     // NSTerm.hh:113-119
        const CFreal mine = (c.viscLaw == CFB200_VISC_SIMPLER_SUTHERLAND) ? c.viscRef*std::pow(T, 1.5)/(c.viscTRef + T) :
                            (c.viscLaw == CFB200_VISC_SUTHERLAND) ? c.viscRef*(c.viscTRef + c.viscSuthConst)/(T + c.viscSuthConst)*std::pow(T/c.viscTRef, 1.5) :
                            c.fixedVisc;
        if (std::abs(mine - ref) > 1e-13*std::abs(ref))
          throw BadValueException(FromHere(), "B200: the viscosity law constants differ from the configured law: set "
                                  "ViscRef / TRef / SuthConst / FixedVisc of this command to the values of NavierStokes.DiffTerm");
      }
    }
  }
  // ---- time integration scalars that the pseudo-steady term reads (FVMCC_PseudoSteadyTimeRhs.cxx:40-46,150-175); the
  // flags of the configured "PseudoSteadyTimeRhs" command are mirrored by this command's options of the same names
  c.timeAnalyticalMatrix = m_timeAnalyticalMatrix ? 1 : 0;
  c.timeUseGlobalDT = m_timeUseGlobalDT ? 1 : 0;
  for (CFuint i = 0; i < nbEqs && i < m_timeZeroDiag.size(); ++i) c.timeZeroDiag[i] = m_timeZeroDiag[i] ? 1 : 0;
  c.fuseTimeRhs = m_fuseTimeRhs ? 1 : 0;
}

// per-call scalars (SURVEY.md section 8b "Scalars crossing per call")
void FVMCC_ComputeRHSB200::refreshStep(cfb200_config& c)
{
  c.resFactor = getResFactor();
  c.computeJacobian = (getMethodData().doComputeJacobian() && !getMethodData().isSysMatrixFrozen()) ? withJacobian() : 0;
  c.iter = SubSystemStatusStack::getActive()->getNbIter();
  c.residual = SubSystemStatusStack::getActive()->getResidual();
  c.cfl = getMethodData().getCFL()->getCFLValue();        // FVMCC_PseudoSteadyTimeRhs.cxx:119
  c.dt = SubSystemStatusStack::getActive()->getDT();      // FVMCC_PseudoSteadyTimeRhs.cxx:147
  // PolyReconstructor options declared DynamicOption (limitIter, limitRes, gradientFactor) and RoeFlux::DiffCoeff may be
  // changed interactively / be functions of the iteration: re-read them
  FVMCC_PolyRec& rec = *getMethodData().getPolyReconstructor();
  c.limitIter = optInt(rec, "limitIter", c.limitIter);
  c.limitRes = optReal(rec, "limitRes", c.limitRes);
  c.gradientFactor = optReal(rec, "gradientFactor", c.gradientFactor);
  c.diffRedCoeff = optReal(*getMethodData().getFluxSplitter(), "DiffCoeff", c.diffRedCoeff);
}

void FVMCC_ComputeRHSB200::setup()
{
  FVMCC_ComputeRHS::setup();
  fillConfig(m_cfg);                   // string matching and option read-back happen once, here
  refreshStep(m_cfg);
  cfb200_config& c = m_cfg;
  int device = m_device;
  if (device < 0) device = PE::GetPE().GetRank("Default") % std::max(1, cfb200_device_count());
  check(cfb200_create(&c, device, &m_h), "cfb200_create");
  uploadMesh();
}

// One cfb200_bc per boundary TRS.  Commands without parameters are recognised by their class; FarFieldEuler2D/3D by
// class + its options.  The Dirichlet-type commands (SuperInlet with a constant Def, SubInletEulerPvtVT,
// SubOutletEulerPvt, NoSlipWallAdiabaticNSPvt, SuperOutlet, ...) all set ghost = 2*value - inner on a subset of the
// variables and copy the others, so mask and value are read off the command itself by applying it to a zero inner
// state on one face of the TRS (ghost = 2*value exactly) and are then verified, bit for bit, against the command on
// the face's real inner state.  Anything that fails the verification is reported as unsupported (type -1).
cfb200_bc FVMCC_ComputeRHSB200::translateBC(FVMCC_BC* bc, GeometricEntity* probe, CFuint nEqs)
{
  cfb200_bc out;
  std::memset(&out, 0, sizeof(out));
  const std::vector<bool>& zg = *bc->getZeroGradientsFlags();
  for (CFuint i = 0; i < nEqs && i < zg.size(); ++i) out.zeroGrad[i] = zg[i] ? 1 : 0;
  const std::string cls = Common::demangle(typeid(*bc).name());
  const auto has = [&](const char* name) { return cls.find(name) != std::string::npos; };
  const auto opt = [&](const char* name) { return std::strtod(bc->getOption(name)->getValueAsString().c_str(), CFNULL); };
  if (has("SuperOutletMHD3DProjection") && !has("Heating")) {       // SuperOutletMHD3DProjection.cxx:34-48: refPhi, coronalBC
    out.type = CFB200_BC_SUPER_OUTLET_MHD;
    const std::string cor = bc->getOption("coronalBC")->getValueAsString();
    out.mask[0] = (cor == "true" || cor == "1") ? 1 : 0;
    out.value[0] = std::strtod(bc->getOption("refPhi")->getValueAsString().c_str(), CFNULL);
    return out;
  }
  if (has("MirrorVelocity")) { out.type = CFB200_BC_MIRROR_VELOCITY; return out; }
  if (has("MirrorEuler2D") || has("MirrorEuler3D") || has("MirrorMHD3DProjection")) {
    out.type = CFB200_BC_MIRROR_EULER_CONS; return out;
  }
  if (has("FarFieldEuler2D") || has("FarFieldEuler3D") || has("FarFieldEulerChar2D") || has("FarFieldEulerChar3D")) {
    // constant far field only (no Def functions, no interpolation file); the option strings carry the values
    // (FarFieldEulerChar2D / 3D derive FarFieldEuler2D / 3D and read the same options)
    const CFuint dim = PhysicalModelStack::getActive()->getDim();
    out.type = has("FarFieldEulerChar") ? CFB200_BC_FARFIELD_EULER_CHAR : CFB200_BC_FARFIELD_EULER;
    out.value[0] = opt("Pressure"); out.value[1] = opt("Uinf"); out.value[2] = opt("Vinf");
    if (dim == 3) out.value[3] = opt("Winf");
    out.value[dim + 1] = opt("Temperature");
    return out;
  }
  if (has("NoSlipWallIsothermalNSvt")) {
    // NoSlipWallIsothermalNSvt.ci:22-83: TWall (dimensional), uWall, TGhostMin; the radiative-equilibrium wall, the
    // fixed-ghost-temperature variant and the "Adiabatic" switch are other commands in disguise: not translated
    const auto flag = [&](const char* name) { return bc->getOption(name)->getValueAsString() == "true"; };
    if (flag("RadEquilibrium") || flag("FixTWallInGhost") || flag("Adiabatic")) { out.type = -1; return out; }
    const CFuint dim = PhysicalModelStack::getActive()->getDim();
    const CFreal tempRef = m_cfg.tempRef;                 // EulerTerm::getTempRef(), read by fillConfig()
    out.type = CFB200_BC_NOSLIP_ISOTHERMAL;
    out.value[0] = opt("TGhostMin")/tempRef;
    std::istringstream uw(bc->getOption("uWall")->getValueAsString());     // "u v [w]", default: all zero
    for (CFuint k = 0; k < dim; ++k) { CFreal v = 0.; uw >> v; out.value[1 + k] = v; }
    out.value[dim + 1] = opt("TWall")/tempRef;
    return out;
  }
  out.type = -1;
  if (probe == CFNULL) { out.type = CFB200_BC_SUPER_OUTLET; return out; }    // empty TRS on this rank
  State& inner = *probe->getState(0);
  State& ghost = *probe->getState(1);
  const RealVector innerBkp = inner, ghostBkp = ghost;
  inner = 0.;
  bc->setGhostState(probe);
  bool any = false;
  for (CFuint i = 0; i < nEqs; ++i) {
    out.mask[i] = (ghost[i] != 0.) ? 1 : 0;
    out.value[i] = 0.5*ghost[i];
    any = any || out.mask[i];
  }
  static_cast<RealVector&>(inner) = innerBkp;
  bc->setGhostState(probe);
  bool linear = true;
  for (CFuint i = 0; i < nEqs; ++i) {
    const CFreal expect = out.mask[i] ? 2.*out.value[i] - inner[i] : inner[i];
    linear = linear && (ghost[i] == expect);
  }
  static_cast<RealVector&>(ghost) = ghostBkp;
  if (linear) out.type = any ? CFB200_BC_DIRICHLET : CFB200_BC_SUPER_OUTLET;
  return out;
}

void FVMCC_ComputeRHSB200::uploadMesh()
{
  // Flattens the sockets and connectivity tables into a cfb200_mesh (include/cfb200.h; field-by-field mapping:
  // INTEGRATION.md §3).  State-indexed arrays hold the cells first, then the ghost states.
  const CFuint dim = PhysicalModelStack::getActive()->getDim();
  DataHandle<State*, GLOBAL> states = socket_states.getDataHandle();
  DataHandle<State*> gstates = socket_gstates.getDataHandle();
  DataHandle<Node*, GLOBAL> nodes = socket_nodes.getDataHandle();
  DataHandle<CFreal> normals = socket_normals.getDataHandle();
  DataHandle<CFreal> faceAreas = socket_faceAreas.getDataHandle();
  DataHandle<CFreal> volumes = socket_volumes.getDataHandle();
  DataHandle<std::vector<State*> > stencil = socket_stencil.getDataHandle();
  const CFuint nCells = states.size(), nGhosts = gstates.size(), nNodes = nodes.size();

  std::vector<CFreal> xNodes(nNodes*dim), xStates((nCells + nGhosts)*dim), vol(nCells);
  std::vector<uint8_t> updatable(nCells);
  for (CFuint i = 0; i < nNodes; ++i)
    for (CFuint k = 0; k < dim; ++k) xNodes[i*dim + k] = (*nodes[i])[k];
  for (CFuint i = 0; i < nCells; ++i) {
    for (CFuint k = 0; k < dim; ++k) xStates[i*dim + k] = states[i]->getCoordinates()[k];
    vol[i] = volumes[i];
    updatable[i] = states[i]->isParUpdatable() ? 1 : 0;
  }
  for (CFuint i = 0; i < nGhosts; ++i)
    for (CFuint k = 0; k < dim; ++k) xStates[(nCells + i)*dim + k] = gstates[i]->getCoordinates()[k];
  // a state pointer -> its index in the state-indexed arrays
  const auto stateIndex = [&](const State* s) -> CFint {
    return s->isGhost() ? (CFint)(nCells + s->getLocalID()) : (CFint)s->getLocalID();
  };


  // Faces in the order of the TRS loop of FVMCC_ComputeRHS::execute (FVMCC_ComputeRHS.cxx:148-300): InnerFaces, then
  // every boundary TRS (FVMCC_MeshDataBuilder.cxx:91-124 creates them in that order).  normals / faceAreas are
  // indexed by face ID in the sockets and by loop position in cfb200_mesh.
  const CFuint nEqs = PhysicalModelStack::getActive()->getNbEq();
  std::vector<CFint> faceL, faceR, faceNodePtr(1, 0), faceNodes, faceBC;
  std::vector<CFreal> nrm, area;
  std::map<CFuint, CFint> loopIndexOfFace;
  std::vector<cfb200_bc> bcs;
  std::map<CFint, std::vector<CFreal> > faceValues;      // per Dirichlet record: value[iFace*nEqs + v]
  std::map<CFint, bool> positionDependent;
  CFint nInner = -1;
  SafePtr<GeometricEntityPool<FaceCellTrsGeoBuilder> > geoBuilder = getMethodData().getFaceCellTrsGeoBuilder();
  FaceCellTrsGeoBuilder::GeoData& geoData = geoBuilder->getDataGE();
  SafePtr<CFMap<CFuint, FVMCC_BC*> > bcMap = getMethodData().getMapBC();
  std::vector<SafePtr<TopologicalRegionSet> > trs = MeshDataStack::getActive()->getTrsList();
  for (CFuint iTRS = 0; iTRS < trs.size(); ++iTRS) {
    SafePtr<TopologicalRegionSet> currTrs = trs[iTRS];
    if (currTrs->getName() == "PartitionFaces" || currTrs->getName() == "InnerCells") continue;
    const bool isBFace = currTrs->hasTag("writable");
    if (!isBFace && nInner >= 0) throw BadValueException(FromHere(), "B200: more than one interior face TRS");
    if (isBFace && nInner < 0) throw BadValueException(FromHere(), "B200: a boundary TRS precedes InnerFaces");
    CFint bcIndex = -1;
    if (isBFace) {
      FVMCC_BC* bc = bcMap->find(iTRS);
      geoData.trs = currTrs; geoData.isBFace = true; geoData.idx = 0;
      GeometricEntity* probe = (currTrs->getLocalNbGeoEnts() > 0) ? geoBuilder->buildGE() : CFNULL;
      bcs.push_back(translateBC(bc, probe, nEqs));
      if (probe != CFNULL) geoBuilder->releaseGE();
      bcIndex = (CFint)bcs.size() - 1;
      if (bcs[bcIndex].type == CFB200_BC_DIRICHLET) m_dirichletTrs.push_back(std::make_pair(iTRS, bcIndex));
    }
    geoData.trs = currTrs;
    geoData.isBFace = isBFace;
    const CFuint nbTrsFaces = currTrs->getLocalNbGeoEnts();
    for (CFuint iFace = 0; iFace < nbTrsFaces; ++iFace) {
      geoData.idx = iFace;
      GeometricEntity* face = geoBuilder->buildGE();
      const CFuint faceID = face->getID();
      loopIndexOfFace[faceID] = (CFint)faceL.size();
      faceL.push_back(stateIndex(face->getState(0)));
      faceR.push_back(stateIndex(face->getState(1)));
      for (CFuint n = 0; n < face->nbNodes(); ++n) faceNodes.push_back((CFint)face->getNode(n)->getLocalID());
      faceNodePtr.push_back((CFint)faceNodes.size());
      for (CFuint k = 0; k < dim; ++k) nrm.push_back(normals[faceID*dim + k]);
      area.push_back(faceAreas[faceID]);
      faceBC.push_back(bcIndex);
      if (isBFace && bcs[bcIndex].type == CFB200_BC_DIRICHLET) {
        // value of the record at this face (a Def function of the position gives a different one per face):
        // the command applied to a zero inner state leaves ghost = 2*value
        State& inner = *face->getState(0);
        State& ghost = *face->getState(1);
        const RealVector innerBkp = inner, ghostBkp = ghost;
        inner = 0.;
        bcMap->find(iTRS)->setGhostState(face);
        for (CFuint v = 0; v < nEqs; ++v) {
          const CFreal value = 0.5*ghost[v];
          faceValues[bcIndex].push_back(value);
          if (bcs[bcIndex].mask[v] && value != bcs[bcIndex].value[v]) positionDependent[bcIndex] = true;
        }
        static_cast<RealVector&>(inner) = innerBkp;
        static_cast<RealVector&>(ghost) = ghostBkp;
      }
      geoBuilder->releaseGE();
    }
    if (!isBFace) nInner = (CFint)faceL.size();
  }


  // cellFaces (the limiter and the residual gather follow the cell's local face order), reconstruction stencil,
  // nodal-extrapolation neighbours and weights (DistanceBasedExtrapolator.ci:105-125)
  SafePtr<ConnectivityTable<CFuint> > cf = MeshDataStack::getActive()->getConnectivity("cellFaces");
  std::vector<CFint> cellFacePtr(1, 0), cellFaces, stencilPtr(1, 0), stencilIdx, nodeNbrPtr(1, 0), nodeNbr;
  std::vector<CFreal> nodeW;
  for (CFuint c = 0; c < nCells; ++c) {
    for (CFuint k = 0; k < cf->nbCols(c); ++k) cellFaces.push_back(loopIndexOfFace[(*cf)(c, k)]);
    cellFacePtr.push_back((CFint)cellFaces.size());
    for (CFuint k = 0; k < stencil[c].size(); ++k) stencilIdx.push_back(stateIndex(stencil[c][k]));
    stencilPtr.push_back((CFint)stencilIdx.size());
  }
  SafePtr<NodalStatesExtrapolator<CellCenterFVMData> > nse = getMethodData().getNodalStatesExtrapolator();
  for (CFuint n = 0; n < nNodes; ++n) {
    const std::vector<State*>& nb = nse->getNodalStateNeighbors(n);
    CFreal sumWeight = 0.0;
    const size_t first = nodeW.size();
    for (CFuint k = 0; k < nb.size(); ++k) {
      nodeNbr.push_back(stateIndex(nb[k]));
      const CFreal weight = 1./MathTools::MathFunctions::getDistance(nb[k]->getCoordinates(), *nodes[n]);
      nodeW.push_back(weight);
      sumWeight += weight;
    }
    for (size_t k = first; k < nodeW.size(); ++k) nodeW[k] /= sumWeight;
    nodeNbrPtr.push_back((CFint)nodeNbr.size());
  }

  cfb200_mesh mesh;
  mesh.dim = dim; mesh.nCells = nCells; mesh.nGhosts = nGhosts; mesh.nNodes = nNodes;
  mesh.nFaces = (CFint)faceL.size(); mesh.nInner = nInner;
  mesh.nodes = &xNodes[0]; mesh.centres = &xStates[0];
  mesh.faceL = &faceL[0]; mesh.faceR = &faceR[0]; mesh.faceNodePtr = &faceNodePtr[0]; mesh.faceNodes = &faceNodes[0];
  mesh.normals = &nrm[0]; mesh.areas = &area[0]; mesh.volumes = &vol[0];
  mesh.cellFacePtr = &cellFacePtr[0]; mesh.cellFaces = &cellFaces[0];
  mesh.stencilPtr = &stencilPtr[0]; mesh.stencil = &stencilIdx[0];
  mesh.nodeNbrPtr = &nodeNbrPtr[0]; mesh.nodeNbr = &nodeNbr[0]; mesh.nodeW = &nodeW[0];
  mesh.faceBC = &faceBC[0]; mesh.updatable = &updatable[0];
  // cell -> nodes in the cell-local order (CorrectedGG2D/3D builds each cell with CellTrsGeoBuilder: the order of the
  // "cellNodes_InnerCells" connectivity, CorrectedDerivativeGG2D.cxx:306-322)
  std::vector<CFint> cellNodePtr(1, 0), cellNodesIdx;
  {
    SafePtr<ConnectivityTable<CFuint> > cn = MeshDataStack::getActive()->getConnectivity("cellNodes_InnerCells");
    for (CFuint c = 0; c < nCells; ++c) {
      for (CFuint k = 0; k < cn->nbCols(c); ++k) cellNodesIdx.push_back((CFint)(*cn)(c, k));
      cellNodePtr.push_back((CFint)cellNodesIdx.size());
    }
  }
  mesh.cellNodePtr = &cellNodePtr[0]; mesh.cellNodes = &cellNodesIdx[0];
  std::vector<CFreal> cvol(faceL.size());
  std::vector<uint8_t> cvSign(faceL.size());
  mesh.cvol = &cvol[0]; mesh.cvSign = &cvSign[0];
  cfb200_mesh_diamond_volumes(&mesh, &cvol[0], &cvSign[0]);
  check(cfb200_set_mesh(m_h, &mesh, (int)bcs.size(), bcs.empty() ? CFNULL : &bcs[0]), "cfb200_set_mesh");
  for (std::map<CFint, bool>::const_iterator it = positionDependent.begin(); it != positionDependent.end(); ++it)
    if (it->second) check(cfb200_set_bc_face_values(m_h, it->first, &faceValues[it->first][0]), "cfb200_set_bc_face_values");

  // NodalExtrapolation = DistanceBasedGMove (DistanceBasedExtrapolatorGMove.cxx:92-118,140-173): the nodes of the TRSs named by
  // `TRSName` that belong to them (socket trsID holds the position of the node's TRS in the extrapolator's ordered list:
  // `TrsPriorityList`, else every boundary TRS in mesh order) get `Values` at `ValuesIdx` after the weighted sum
  if (hasOpt(*nse, "ValuesIdx") && !optStr(*nse, "ValuesIdx", "").empty()) {
    if (optStr(*nse, "RadEquilibrium", "0") == "1" || optStr(*nse, "RadEquilibrium", "false") == "true" ||
        !optStr(*nse, "FileNameTw", "").empty() || optInt(*nse, "NbIterAdiabatic", 0) > 0)
      throw Common::NotImplementedException(FromHere(), "B200: DistanceBasedGMove with RadEquilibrium / file-interpolated nodal values / NbIterAdiabatic");
    if (!socket_trsID.isConnected()) throw Common::NoSuchStorageException(FromHere(), "B200: socket trsID is needed by DistanceBasedGMove");
    DataHandle<CFint> trsID = socket_trsID.getDataHandle();
    std::vector<std::string> ordered, wanted;
    { std::istringstream in(optStr(*nse, "TrsPriorityList", "")); std::string w; while (in >> w) ordered.push_back(w); }
    if (ordered.empty()) {
      std::vector<SafePtr<TopologicalRegionSet> > trsList = MeshDataStack::getActive()->getTrsList();
      for (CFuint i = 0; i < trsList.size(); ++i) {
        const std::string nm = trsList[i]->getName();
        if (nm != "InnerCells" && nm != "InnerFaces" && nm != "PartitionFaces") ordered.push_back(nm);
      }
    }
    { std::istringstream in(optStr(*nse, "TRSName", "")); std::string w; while (in >> w) wanted.push_back(w); }
    std::vector<uint8_t> nodeFlag(nNodes, 0);
    for (CFuint n = 0; n < nNodes; ++n) {
      const CFint id = trsID[n];
      if (id >= 0 && (size_t)id < ordered.size() && std::find(wanted.begin(), wanted.end(), ordered[id]) != wanted.end()) nodeFlag[n] = 1;
    }
    std::vector<CFint> idx; std::vector<CFreal> val;
    { std::istringstream in(optStr(*nse, "ValuesIdx", "")); CFint v; while (in >> v) idx.push_back(v); }
    { std::istringstream in(optStr(*nse, "Values", "")); CFreal v; while (in >> v) val.push_back(v); }
    if (idx.size() != val.size()) throw Common::BadValueException(FromHere(), "B200: DistanceBasedGMove ValuesIdx / Values sizes differ");
    check(cfb200_set_nodal_prescribed(m_h, &nodeFlag[0], (CFint)idx.size(), &idx[0], &val[0]), "cfb200_set_nodal_prescribed");
  }
}

// A SuperInlet whose Def depends on time or on the iteration number (SuperInlet.cxx:156-209 evaluates it inside every
// setGhostState) changes its values between evaluations: with RefreshInletValues the command is applied again to a zero
// inner state on every face of its TRS (ghost = 2*value) and the values are re-sent (cfb200_set_bc_face_values).
void FVMCC_ComputeRHSB200::refreshDirichletValues()
{
  const CFuint nEqs = PhysicalModelStack::getActive()->getNbEq();
  std::vector<SafePtr<TopologicalRegionSet> > trsList = MeshDataStack::getActive()->getTrsList();
  SafePtr<GeometricEntityPool<FaceCellTrsGeoBuilder> > geoBuilder = getMethodData().getFaceCellTrsGeoBuilder();
  FaceCellTrsGeoBuilder::GeoData& geoData = geoBuilder->getDataGE();
  SafePtr<CFMap<CFuint, FVMCC_BC*> > bcMap = getMethodData().getMapBC();
  std::vector<CFreal> values;
  for (CFuint k = 0; k < m_dirichletTrs.size(); ++k) {
    const CFuint iTRS = m_dirichletTrs[k].first;
    SafePtr<TopologicalRegionSet> currTrs = trsList[iTRS];
    geoData.trs = currTrs; geoData.isBFace = true;
    const CFuint nbTrsFaces = currTrs->getLocalNbGeoEnts();
    if (nbTrsFaces == 0) continue;
    values.assign((size_t)nbTrsFaces*nEqs, 0.);
    for (CFuint iFace = 0; iFace < nbTrsFaces; ++iFace) {
      geoData.idx = iFace;
      GeometricEntity* face = geoBuilder->buildGE();
      State& inner = *face->getState(0);
      State& ghost = *face->getState(1);
      const RealVector innerBkp = inner, ghostBkp = ghost;
      inner = 0.;
      bcMap->find(iTRS)->setGhostState(face);
      for (CFuint v = 0; v < nEqs; ++v) values[(size_t)iFace*nEqs + v] = 0.5*ghost[v];
      static_cast<RealVector&>(inner) = innerBkp;
      static_cast<RealVector&>(ghost) = ghostBkp;
      geoBuilder->releaseGE();
    }
    check(cfb200_set_bc_face_values(m_h, m_dirichletTrs[k].second, &values[0]), "cfb200_set_bc_face_values");
  }
}

void FVMCC_ComputeRHSB200::execute()
{
  DataHandle<State*, GLOBAL> states = socket_states.getDataHandle();
  DataHandle<CFreal> rhs = socket_rhs.getDataHandle();
  DataHandle<CFreal> updateCoeff = socket_updateCoeff.getDataHandle();
  refreshStep(m_cfg);
  check(cfb200_set_step(m_h, &m_cfg), "cfb200_set_step");
  if (m_refreshInletValues) refreshDirichletValues();
  // states are contiguous in the global array (FVMCC_ComputeRHSCell.cu:743)
  check(cfb200_set_states(m_h, states.getGlobalArray()->ptr(), states.size()), "cfb200_set_states");
  check(cfb200_set_update_coeff(m_h, &updateCoeff[0]), "cfb200_set_update_coeff");
  if (m_cfg.fuseTimeRhs) {
    // pastStates socket of the time integrator (FVMCC_PseudoSteadyTimeRhs.cxx:199-207): contiguous like the states
    if (!socket_pastStates.isConnected())
      throw BadValueException(FromHere(), "B200: FuseTimeRhs needs the pastStates socket of an implicit time integrator");
    DataHandle<State*> pastStates = socket_pastStates.getDataHandle();
    m_past.resize(states.size()*m_cfg.nEqs);
    for (CFuint i = 0; i < states.size(); ++i)
      for (CFint v = 0; v < m_cfg.nEqs; ++v) m_past[i*m_cfg.nEqs + v] = (*pastStates[i])[v];
    check(cfb200_set_past_states(m_h, &m_past[0], states.size()), "cfb200_set_past_states");
  }
  // ghost states are rebuilt on the device by the same boundary conditions (applyBC = 1)
  check(cfb200_compute_rhs(m_h, 1, &rhs[0], &updateCoeff[0]), "cfb200_compute_rhs");
  if (m_cfg.computeJacobian) insertJacobian();
}

// The block Jacobian -> the LSS matrix, as FVMCC_ComputeRhsJacob does it (FVMCC_ComputeRhsJacob.cxx:126-176: fill a
// BlockAccumulator, LSSMatrix::addValues(acc), LSSMatrix.hh:156-162).  One accumulator of 1 x n blocks per block ROW of
// an updatable cell (n = 1 + its interior faces); rows of non-updatable cells are not touched (the reference never adds
// to them, FVMCC_ComputeRhsJacob.cxx:108-119; FVMCC_ComputeRhsJacobCell.ci:281,296-299 passes row index -1 for them).
// BlockAccumulator::setValue(i, j, ib, jb, v) addresses entry [ib][jb] of sub-block (i, j) whatever index coefficients
// the LSS installed (BlockAccumulator.hh:289-296: PETSc row-oriented blocked layout by default) -- the same [ib][jb] the
// BSR blocks hold.  setRowIndex / setColIndex translate local state IDs through the LSSIdxMapping (:183-207).
// Device-resident consumers (PETSc MATMPIAIJCUSPARSE, Paralution) take cfb200_bsr_view / cfb200_export_csr_device instead
// and skip this host copy (HostJacobian = false).
void FVMCC_ComputeRHSB200::insertJacobian()
{
  SafePtr<LinearSystemSolver> lss = getMethodData().getLinearSystemSolver()[0];
  SafePtr<LSSMatrix> jacobMatrix = lss->getMatrix();
  const CFuint nbEqs = PhysicalModelStack::getActive()->getNbEq();
  DataHandle<State*, GLOBAL> states = socket_states.getDataHandle();
  const CFuint nbCells = states.size();
  if (m_rowPtr.empty()) {
    m_rowPtr.resize(nbCells + 1); m_colInd.resize(cfb200_bsr_nnzb(m_h));
    check(cfb200_bsr_pattern(m_h, &m_rowPtr[0], &m_colInd[0]), "cfb200_bsr_pattern");
    m_bsr.resize(m_colInd.size()*nbEqs*nbEqs);
  }
  // FVMCC_ComputeRhsJacob::initializeComputationRHS resets the matrix itself (FVMCC_ComputeRhsJacob.cxx:413-416); this
  // command stands in its place
  jacobMatrix->resetToZeroEntries();
  if (!m_hostJacobian) return;          // the device view is consumed by a device-side LSS
  check(cfb200_get_bsr_values(m_h, &m_bsr[0]), "cfb200_get_bsr_values");
  for (CFuint iCell = 0; iCell < nbCells; ++iCell) {
    if (!states[iCell]->isParUpdatable()) continue;
    const CFint first = m_rowPtr[iCell], nCols = m_rowPtr[iCell + 1] - first;
    BlockAccumulator*& acc = m_rowAcc[nCols];
    if (acc == CFNULL) acc = lss->createBlockAccumulator(1, nCols, nbEqs);     // LinearSystemSolver.hh:70
    acc->setRowIndex(0, (CFint)iCell);
    for (CFint j = 0; j < nCols; ++j) {
      acc->setColIndex(j, m_colInd[first + j]);
      const CFreal* blk = &m_bsr[(size_t)(first + j)*nbEqs*nbEqs];
      for (CFuint ib = 0; ib < nbEqs; ++ib)
        for (CFuint jb = 0; jb < nbEqs; ++jb) acc->setValue(0, j, ib, jb, blk[ib*nbEqs + jb]);
    }
    jacobMatrix->addValues(*acc);
  }
}

void FVMCC_ComputeRHSB200::unsetup()
{
  if (m_h != CFNULL) { cfb200_destroy(m_h); m_h = CFNULL; }
  for (std::map<CFint, BlockAccumulator*>::iterator it = m_rowAcc.begin(); it != m_rowAcc.end(); ++it) delete it->second;
  m_rowAcc.clear();
  FVMCC_ComputeRHS::unsetup();
}

} } }
