// See FVMCC_ComputeRHSB200.hh.  Written against the COOLFluiD API of /root/reference; not compiled here.
#include <cstdlib>
#include <cstring>
#include <map>
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
#include "NavierStokes/EulerTerm.hh"
#include "Framework/NodalStatesExtrapolator.hh"
#include "Framework/ConnectivityTable.hh"
#include "MathTools/MathFunctions.hh"
#include "Common/DemangledTypeID.hh"
#include "FiniteVolume/FVMCC_BC.hh"

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
  return result;
}

void FVMCC_ComputeRHSB200::defineConfigOptions(Config::OptionList& options)
{
  options.addConfigOption< CFint >("Device", "CUDA device ordinal (-1: rank % deviceCount).");
  options.addConfigOption< bool >("HostJacobian", "Copy the block Jacobian to the host LSS matrix (CPU PETSc).");
}

FVMCC_ComputeRHSB200::FVMCC_ComputeRHSB200(const std::string& name) :
  FVMCC_ComputeRHS(name), socket_stencil("stencil"), m_h(CFNULL)
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
  if (has("MirrorVelocity")) { out.type = CFB200_BC_MIRROR_VELOCITY; return out; }
  if (has("MirrorEuler2D") || has("MirrorEuler3D") || has("MirrorMHD3DProjection")) {
    out.type = CFB200_BC_MIRROR_EULER_CONS; return out;
  }
  if (has("FarFieldEuler2D") || has("FarFieldEuler3D")) {
    // constant far field only (no Def functions, no interpolation file); the option strings carry the values
    const CFuint dim = PhysicalModelStack::getActive()->getDim();
    out.type = CFB200_BC_FARFIELD_EULER;
    out.value[0] = opt("Pressure"); out.value[1] = opt("Uinf"); out.value[2] = opt("Vinf");
    if (dim == 3) out.value[3] = opt("Winf");
    out.value[dim + 1] = opt("Temperature");
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
  std::vector<CFreal> cvol(faceL.size());
  std::vector<uint8_t> cvSign(faceL.size());
  mesh.cvol = &cvol[0]; mesh.cvSign = &cvSign[0];
  cfb200_mesh_diamond_volumes(&mesh, &cvol[0], &cvSign[0]);
  check(cfb200_set_mesh(m_h, &mesh, (int)bcs.size(), bcs.empty() ? CFNULL : &bcs[0]), "cfb200_set_mesh");
  for (std::map<CFint, bool>::const_iterator it = positionDependent.begin(); it != positionDependent.end(); ++it)
    if (it->second) check(cfb200_set_bc_face_values(m_h, it->first, &faceValues[it->first][0]), "cfb200_set_bc_face_values");
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
