// cfb200 — host side of the engine: owns the device-resident mesh/state/Jacobian storage, builds the
// gather tables once (cell->faces in loop order, LSQ edge order, BSR slots), and sequences the kernels
// of one ComputeRHS (+Jacobian) evaluation.  Mirrors the command interface of the reference
// (setup / execute / unsetup of FVMCC_ComputeRHS, plugins/FiniteVolume/FVMCC_ComputeRHS.hh:37-77).
//
// `BK` is the execution backend.  The product instantiates it ONLY with the CUDA backend
// (cfb_capi.cu); tests/hostsim instantiates it with a host loop backend to check the kernel
// functors against the oracle on machines without a GPU.  That host instantiation is test
// infrastructure and is never linked into libcfb200.so.
#pragma once
#include <vector>
#include <string>
#include <cstring>
#include <algorithm>
#include <stdexcept>
#include <functional>
#include <cstdlib>
#include <cmath>
#include "cfb_kernels.cuh"

namespace cfb {

// Roe; AUSM+ and AUSM+up with the perfect-gas interface sound speeds (choiceA12 4/5 are the TCNEQ variants).
inline bool splitterSupported(const cfb200_config& c) {
  if (c.fuseTimeRhs && c.timeUseGlobalDT) return false;   // the global minimum needs the complete updateCoeff first
  // "NumJacobAnalytic" (a19): the analytic flux Jacobian exists for AUSM+up, in conservative variables (inviscid) and in
  // Puvt / Pvt; the analytic diffusive Jacobian (NSFluxAndJacob) exists for the Puvt / Pvt variable sets with the
  // DiamondVolume computers and Corrected2D (CorrectedDerivative3D has no getGradientsJacob)
  if (c.computeJacobian == 2) {
    if (!(c.physicalModel == CFB200_PM_EULER && c.fluxSplitter == CFB200_FS_AUSMPLUSUP && c.machInf > 0.)) return false;
    if (c.diffusive != CFB200_DIFF_NONE && c.updateVar != CFB200_UV_PVT) return false;
    if (c.diffusive != CFB200_DIFF_NONE && c.derivativeStrategy == CFB200_DV_CORRECTED && c.dim != 2) return false;
    if (c.diffusive != CFB200_DIFF_NONE && (c.derivativeStrategy == CFB200_DV_STATEDIFF ||
                                            c.derivativeStrategy == CFB200_DV_CORRECTED_GG)) return false;   // getGradientsJacob: stubs
  }
  // Corrected2D/3D reads the cell gradients (uX uY uZ sockets), which only the linear reconstruction computes
  if (c.diffusive == CFB200_DIFF_NS && c.derivativeStrategy != CFB200_DV_DIAMOND &&
      c.derivativeStrategy != CFB200_DV_STATEDIFF && c.derivativeStrategy != CFB200_DV_CORRECTED_GG &&
      !(c.derivativeStrategy == CFB200_DV_CORRECTED && c.polyRec != CFB200_PR_CONSTANT))
    return false;
  // Data.SourceTerm: none, or MHDConsACAST (its phi damping term) on MHD3DProjection, FD Jacobian (useAnalyticalJacob is off there)
  if (c.sourceTerm != CFB200_ST_NONE &&
      !(c.sourceTerm == CFB200_ST_MHD_CONS_ACA && c.physicalModel == CFB200_PM_MHD3D_PROJECTION && c.computeJacobian != 2 &&
        !(c.mhdDissipCoeff < 0.))) return false;
  // background potential field: none or Dipole, MHD3DProjection only ("PFSS" needs the reference's harmonics data)
  if (c.mhdPotentialBType != 0 && !(c.mhdPotentialBType == 1 && c.physicalModel == CFB200_PM_MHD3D_PROJECTION)) return false;
  if (c.physicalModel == CFB200_PM_MHD3D_PROJECTION)      // MHD: centred splitters, conservative variables, inviscid
    return (c.fluxSplitter == CFB200_FS_LAXFRIED || c.fluxSplitter == CFB200_FS_HLL || c.fluxSplitter == CFB200_FS_CENTRED) &&
           c.updateVar == CFB200_UV_CONS &&
           c.diffusive == CFB200_DIFF_NONE;
  if (c.fluxSplitter == CFB200_FS_ROE) return c.entropyFixID >= 0 && c.entropyFixID <= 2;
  if (c.fluxSplitter == CFB200_FS_ROESA) return c.entropyFixID >= 0 && c.entropyFixID <= 3;
  if (c.fluxSplitter == CFB200_FS_AUSMPLUS || c.fluxSplitter == CFB200_FS_AUSMPLUSUP)
    return c.choiceA12 >= 1 && c.choiceA12 <= 3 && (c.fluxSplitter == CFB200_FS_AUSMPLUS || c.machInf > 0.);
  if (c.fluxSplitter == CFB200_FS_LAXFRIED || c.fluxSplitter == CFB200_FS_HLL || c.fluxSplitter == CFB200_FS_CENTRED) return true;
  return false;
}
// kernel family of the splitter: 0 Roe, 1 AUSM+ / AUSM+up, 2 Lax-Friedrichs / HLL, 3 Roe with the entropy fix, 4 RoeSA
inline int splitterFamily(const cfb200_config& c) {
  if (c.fluxSplitter == CFB200_FS_ROE) return (c.entropyFixID != 0) ? 3 : 0;
  if (c.fluxSplitter == CFB200_FS_ROESA) return 4;
  return (c.fluxSplitter == CFB200_FS_LAXFRIED || c.fluxSplitter == CFB200_FS_HLL || c.fluxSplitter == CFB200_FS_CENTRED) ? 2 : 1;
}

struct DeviceError : std::runtime_error {
  int code;
  DeviceError(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};

template <class BK>
class Engine {
public:
  Cfg cfg;
  Dev d;
  BK bk;
  std::string err;
  int nBC = 0;
  int64_t nnzb = 0;
  int nNonUpd = 0;
  bool meshSet = false, statesSet = false;
  bool timeTermApplied = false;     // the last computeRhs applied the pseudo-steady time term itself (cfg.fuseTimeRhs)
  int lastStatus = 0;
  uint8_t* nodePresc = nullptr; int* nodePrescIdx = nullptr; double* nodePrescVal = nullptr;   // behind Dev::nodePresc*
  double* bcFaceVal = nullptr; uint8_t* bcFaceFlag = nullptr;      // device copies behind Dev::bcFaceVal / bcFaceFlag
  std::vector<double> hBcFaceVal; std::vector<uint8_t> hBcFaceFlag; std::vector<int32_t> hGhostBC;
  // FD loop of the Roe families (CFB200_ROE_PIPELINE, for A/B measurements): 0 = rolled loop, 1 = software-pipelined,
  // 2 = software-pipelined in two phases with three resident blocks per SM (DiamondVolume; Corrected uses 1)
  int roePipeline = 1;
  int splitD2H = 1;         // host copies requested: face kernel + residual gather in two parts (CFB200_SPLIT_D2H=0: one)
  int splitFace = 0, splitCells = 0;   // interior faces [0, splitFace) complete the cells [0, splitCells)
  int faceShape = 2;       // resident 128-thread blocks per SM the interior-face Jacobian kernel is compiled for
                           // (CudaBackend::launchStaged): 2 -> 255 registers, no spills: measured fastest on B200 (profiles/)
  // multi-GPU (cfb_halo.cuh): a state halo exchange may be in flight when computeRhs is called; the first nIndep
  // cells have no received cell in their reconstruction stencil, so their gradients overlap with the exchange
  int nIndep = 0;
  // the same set given as a list (cfb200_set_independent_cells): local cell numbering then stays free, e.g. ascending
  // global ID, which keeps every gather order of the single-domain run; depCells = the complement
  int* indepCells = nullptr; int* depCells = nullptr; int nIndepList = 0;
  bool haloPending = false;
  std::function<void()> haloWait;
  // every entry point that reads or writes U outside computeRhs joins an exchange that is still in flight first
  void joinHalo() {
    if (!haloPending) return;
    haloPending = false;
    if (haloWait) haloWait();
  }
  std::vector<void*> owned;
  void* solverExt = nullptr; void (*solverExtFree)(void*) = nullptr;     // scratch of the device linear solver (cfb_capi.cu)
  int* nonUpdCells = nullptr;
  std::vector<int32_t> hRowPtr, hColInd;

  explicit Engine(const Cfg& c, int device) : cfg(c), bk(device) {
    std::memset(&d, 0, sizeof(d));
    if (const char* v = std::getenv("CFB200_ROE_PIPELINE")) roePipeline = std::atoi(v);
    if (const char* v = std::getenv("CFB200_SPLIT_D2H")) splitD2H = std::atoi(v);
  }
  ~Engine() { if (solverExt && solverExtFree) solverExtFree(solverExt); for (void* p : owned) bk.free(p); }

  template <class T> T* dalloc(size_t n) {
    T* p = (T*)bk.alloc(std::max<size_t>(n, 1)*sizeof(T));
    owned.push_back(p);
    return p;
  }
  template <class T> T* upload(const std::vector<T>& v) {
    T* p = dalloc<T>(v.size());
    if (!v.empty()) bk.h2d(p, v.data(), v.size()*sizeof(T));
    return p;
  }
  template <class T> T* upload(const T* src, size_t n) {
    T* p = dalloc<T>(n);
    if (n) bk.h2d(p, src, n*sizeof(T));
    return p;
  }
  static std::vector<double> toSoA(const double* a, size_t n, int k) {
    std::vector<double> r(n*k);
    for (size_t i = 0; i < n; ++i) for (int j = 0; j < k; ++j) r[(size_t)j*n + i] = a[i*k + j];
    return r;
  }

  int q() const { return cfg.nEqs; }

  // ---- storage permutation (cfg.renumber) -----------------------------------------------------------
  // Cells along the Morton curve of their centres, interior faces by (lower, higher) storage cell; boundary faces, ghosts
  // and nodes keep their places.  The permuted copy of the caller's mesh keeps every LIST ORDER of the original (a cell's
  // faces, its stencil, a node's neighbours): only the IDs inside change.  cellRef / faceRef give the caller's IDs back.
  struct Permuted {
    std::vector<double> centres, normals, areas, volumes, cvol;
    std::vector<int32_t> faceL, faceR, faceNodePtr, faceNodes, cellFacePtr, cellFaces, stencilPtr, stencil, nodeNbr, faceBC;
    std::vector<int32_t> cellNodePtr, cellNodes;
    std::vector<uint8_t> cvSign, updatable;
    cfb200_mesh view;
  };
  std::vector<int32_t> hCellRef, hCellInv, hFaceRef;     // storage -> caller, caller -> storage, storage face -> caller face
  int* dStateRef = nullptr; int* dFaceRef = nullptr;     // device: [nStates] (ghosts map to themselves), [nFaces]
  uint8_t* dUpdatableByRow = nullptr;                    // updatable flag by BSR row (= caller's cell ID)
  static uint64_t spread3(uint64_t x) {                  // 21 bits -> every third bit
    x &= 0x1fffff; x = (x | x << 32) & 0x1f00000000ffffULL; x = (x | x << 16) & 0x1f0000ff0000ffULL;
    x = (x | x << 8) & 0x100f00f00f00f00fULL; x = (x | x << 4) & 0x10c30c30c30c30c3ULL; x = (x | x << 2) & 0x1249249249249249ULL;
    return x;
  }
  void buildPermutation(const cfb200_mesh& m, Permuted& P) {
    const int D = m.dim, nc = m.nCells, ng = m.nGhosts, nf = m.nFaces, ni = m.nInner, nn = m.nNodes;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int c = 0; c < nc; ++c) for (int k = 0; k < D; ++k) { lo[k] = std::min(lo[k], m.centres[(size_t)c*D + k]); hi[k] = std::max(hi[k], m.centres[(size_t)c*D + k]); }
    std::vector<std::pair<uint64_t, int32_t> > key(nc);
    for (int c = 0; c < nc; ++c) {
      uint64_t code = 0;
      for (int k = 0; k < D; ++k) {
        const double t = (hi[k] > lo[k]) ? (m.centres[(size_t)c*D + k] - lo[k])/(hi[k] - lo[k]) : 0.;
        code |= spread3((uint64_t)(t*2097151.0)) << k;
      }
      key[c] = std::make_pair(code, (int32_t)c);
    }
    std::sort(key.begin(), key.end());
    hCellRef.resize(nc); hCellInv.resize(nc);
    for (int i = 0; i < nc; ++i) { hCellRef[i] = key[i].second; hCellInv[key[i].second] = i; }
    auto st = [&](int s) { return s < nc ? hCellInv[s] : s; };      // state id: cells move, ghosts stay
    // interior faces by their storage cells
    std::vector<std::pair<std::pair<int32_t, int32_t>, int32_t> > fk(ni);
    for (int f = 0; f < ni; ++f) {
      const int a = hCellInv[m.faceL[f]], b = hCellInv[m.faceR[f]];
      fk[f] = std::make_pair(std::make_pair(std::min(a, b), std::max(a, b)), (int32_t)f);
    }
    std::sort(fk.begin(), fk.end());
    hFaceRef.resize(nf);
    std::vector<int32_t> finv(nf);
    for (int f = 0; f < ni; ++f) hFaceRef[f] = fk[f].second;
    for (int f = ni; f < nf; ++f) hFaceRef[f] = f;
    for (int f = 0; f < nf; ++f) finv[hFaceRef[f]] = f;
    const int ns = nc + ng;
    P.centres.resize((size_t)ns*D);
    for (int i = 0; i < ns; ++i) { const int o = i < nc ? hCellRef[i] : i; for (int k = 0; k < D; ++k) P.centres[(size_t)i*D + k] = m.centres[(size_t)o*D + k]; }
    P.volumes.resize(nc); P.updatable.resize(nc);
    for (int i = 0; i < nc; ++i) { P.volumes[i] = m.volumes[hCellRef[i]]; P.updatable[i] = m.updatable[hCellRef[i]]; }
    P.faceL.resize(nf); P.faceR.resize(nf); P.normals.resize((size_t)nf*D); P.areas.resize(nf); P.cvol.resize(nf);
    P.cvSign.resize(nf); P.faceBC.resize(nf); P.faceNodePtr.assign(1, 0);
    for (int f = 0; f < nf; ++f) {
      const int o = hFaceRef[f];
      P.faceL[f] = st(m.faceL[o]); P.faceR[f] = st(m.faceR[o]);       // the orientation (which cell is L) is the caller's
      for (int k = 0; k < D; ++k) P.normals[(size_t)f*D + k] = m.normals[(size_t)o*D + k];
      P.areas[f] = m.areas[o]; P.cvol[f] = m.cvol[o]; P.cvSign[f] = m.cvSign[o]; P.faceBC[f] = m.faceBC[o];
      P.faceNodes.insert(P.faceNodes.end(), m.faceNodes + m.faceNodePtr[o], m.faceNodes + m.faceNodePtr[o+1]);
      P.faceNodePtr.push_back((int32_t)P.faceNodes.size());
    }
    P.cellFacePtr.assign(1, 0); P.stencilPtr.assign(1, 0);
    for (int i = 0; i < nc; ++i) {
      const int o = hCellRef[i];
      for (int k = m.cellFacePtr[o]; k < m.cellFacePtr[o+1]; ++k) P.cellFaces.push_back(finv[m.cellFaces[k]]);
      P.cellFacePtr.push_back((int32_t)P.cellFaces.size());
      for (int k = m.stencilPtr[o]; k < m.stencilPtr[o+1]; ++k) P.stencil.push_back(st(m.stencil[k]));
      P.stencilPtr.push_back((int32_t)P.stencil.size());
    }
    P.nodeNbr.resize((size_t)m.nodeNbrPtr[nn]);
    for (size_t k = 0; k < P.nodeNbr.size(); ++k) P.nodeNbr[k] = st(m.nodeNbr[k]);
    P.view = m;
    if (m.cellNodePtr && m.cellNodes) {
      P.cellNodePtr.assign(1, 0);
      for (int i = 0; i < nc; ++i) {
        const int o = hCellRef[i];
        P.cellNodes.insert(P.cellNodes.end(), m.cellNodes + m.cellNodePtr[o], m.cellNodes + m.cellNodePtr[o+1]);
        P.cellNodePtr.push_back((int32_t)P.cellNodes.size());
      }
      P.view.cellNodePtr = P.cellNodePtr.data(); P.view.cellNodes = P.cellNodes.data();
    }
    P.view.centres = P.centres.data(); P.view.faceL = P.faceL.data(); P.view.faceR = P.faceR.data();
    P.view.faceNodePtr = P.faceNodePtr.data(); P.view.faceNodes = P.faceNodes.data(); P.view.normals = P.normals.data();
    P.view.areas = P.areas.data(); P.view.volumes = P.volumes.data(); P.view.cellFacePtr = P.cellFacePtr.data();
    P.view.cellFaces = P.cellFaces.data(); P.view.stencilPtr = P.stencilPtr.data(); P.view.stencil = P.stencil.data();
    P.view.nodeNbr = P.nodeNbr.data(); P.view.cvol = P.cvol.data(); P.view.cvSign = P.cvSign.data();
    P.view.faceBC = P.faceBC.data(); P.view.updatable = P.updatable.data();
  }
  // the caller's ID of a storage cell / face (the identity without renumbering): what defines every summation order
  int cref(int c) const { return hCellRef.empty() ? c : hCellRef[c]; }
  int fref(int f) const { return hFaceRef.empty() ? f : hFaceRef[f]; }
  int cinv(int c) const { return hCellInv.empty() ? c : hCellInv[c]; }

  // ---- setup: static data + gather tables --------------------------------------------------------
  void setMesh(const cfb200_mesh& mIn, int nbc, const cfb200_bc* bcs) {
    if (mIn.dim != cfg.dim) throw std::invalid_argument("mesh dimension differs from the configuration");
    Permuted perm;
    if (cfg.renumber) buildPermutation(mIn, perm);
    const cfb200_mesh& m = cfg.renumber ? perm.view : mIn;
    const int D = m.dim, nc = m.nCells, ng = m.nGhosts, nf = m.nFaces, nn = m.nNodes, ni = m.nInner;
    const int ns = nc + ng;
    const int Q = q();
    d.dim = D; d.nCells = nc; d.nGhosts = ng; d.nStates = ns; d.nNodes = nn; d.nFaces = nf; d.nInner = ni;
    nBC = nbc;
    d.nodes = upload(toSoA(m.nodes, nn, D));
    d.centres = upload(toSoA(m.centres, ns, D));
    d.normals = upload(toSoA(m.normals, nf, D));
    d.faceL = upload(m.faceL, nf); d.faceR = upload(m.faceR, nf);
    d.faceNodePtr = upload(m.faceNodePtr, (size_t)nf + 1);
    d.faceNodes = upload(m.faceNodes, (size_t)m.faceNodePtr[nf]);
    d.areas = upload(m.areas, nf);
    d.volumes = upload(m.volumes, nc);
    d.cellFacePtr = upload(m.cellFacePtr, (size_t)nc + 1);
    const size_t ncf = m.cellFacePtr[nc];
    d.cellFaces = upload(m.cellFaces, ncf);
    d.stencilPtr = upload(m.stencilPtr, (size_t)nc + 1);
    d.stencil = upload(m.stencil, (size_t)m.stencilPtr[nc]);
    d.nodeNbrPtr = upload(m.nodeNbrPtr, (size_t)nn + 1);
    d.nodeNbr = upload(m.nodeNbr, (size_t)m.nodeNbrPtr[nn]);
    d.nodeW = upload(m.nodeW, (size_t)m.nodeNbrPtr[nn]);
    d.cvol = upload(m.cvol, nf);
    d.cvSign = upload(m.cvSign, nf);
    d.bndStPtr = nullptr; d.bndSt = nullptr;
    if (cfg.derivativeStrategy == CFB200_DV_CORRECTED && cfg.diffusive == CFB200_DIFF_NS) {
      // gradient stencil of each boundary face (CorrectedDerivative2D.cxx:376-414): the left cell, then, in the order of
      // the left cell's reconstruction stencil, the ghosts other than the face's own and the cells touching a face node
      std::vector<int32_t> sp(1, 0), st;
      for (int g = 0; g < ng; ++g) {
        const int f = ni + g, L = m.faceL[f], R = m.faceR[f];
        const int32_t* fnB = m.faceNodes + m.faceNodePtr[f];
        const int32_t* fnE = m.faceNodes + m.faceNodePtr[f+1];
        st.push_back(L);
        for (int k = m.stencilPtr[L]; k < m.stencilPtr[L+1]; ++k) {
          const int s = m.stencil[k];
          bool member = (s >= nc) ? (s != R) : false;
          if (s < nc)
            for (const int32_t* pn = fnB; pn != fnE && !member; ++pn) {
              const int32_t* nb = m.nodeNbr + m.nodeNbrPtr[*pn];
              const int32_t* ne = m.nodeNbr + m.nodeNbrPtr[*pn + 1];
              member = std::find(nb, ne, s) != ne;
            }
          if (member) st.push_back(s);
        }
        sp.push_back((int32_t)st.size());
      }
      d.bndStPtr = upload(sp); d.bndSt = upload(st);
    }
    d.ggPtr = nullptr; d.ggNode = nullptr; d.ggAv = nullptr;
    if (cfg.derivativeStrategy == CFB200_DV_CORRECTED_GG && cfg.diffusive == CFB200_DIFF_NS) {
      // CorrectedGG2D/3D (CorrectedDerivativeGG2D.cxx:300-397,437-476): (i) per cell the nodes of its faces in the order of
      // computeGradientsGG (local face by local face, the cell-local face-node table); (ii) per face the entries of the
      // control-volume list its "average state" reads: the one FOLLOWING each face node of the L cell (see the oracle);
      // (iii) per boundary face the states across the other faces of the L cell that share a node with it.
      if (!m.cellNodePtr || !m.cellNodes) throw std::invalid_argument("CorrectedGG needs cfb200_mesh::cellNodePtr / cellNodes");
      static const int tri[3][2] = {{0,1},{1,2},{2,0}};
      static const int quad[4][2] = {{0,1},{1,2},{2,3},{3,0}};
      static const int tet[4][3] = {{0,1,2},{0,3,1},{1,3,2},{0,2,3}};
      static const int hex[6][4] = {{0,3,2,1},{4,5,6,7},{0,1,5,4},{1,2,6,5},{2,3,7,6},{0,4,7,3}};
      std::vector<int32_t> gp(1, 0), gn;
      for (int c = 0; c < nc; ++c) {
        const int32_t* cn = m.cellNodes + m.cellNodePtr[c];
        const int ncn = m.cellNodePtr[c+1] - m.cellNodePtr[c];
        const int nfc = m.cellFacePtr[c+1] - m.cellFacePtr[c];
        const bool ok = (D == 2) ? ((ncn == 3 && nfc == 3) || (ncn == 4 && nfc == 4)) : ((ncn == 4 && nfc == 4) || (ncn == 8 && nfc == 6));
        if (!ok) throw std::invalid_argument("CorrectedGG: triangles, quadrilaterals, tetrahedra and hexahedra only");
        for (int iFace = 0; iFace < nfc; ++iFace) {
          const int nbN = (D == 2) ? 2 : (ncn == 4 ? 3 : 4);
          const int* dof = (D == 2) ? (ncn == 3 ? tri[iFace] : quad[iFace]) : (ncn == 4 ? tet[iFace] : hex[iFace]);
          for (int j = 0; j < nbN; ++j) gn.push_back(cn[dof[j]]);
        }
        gp.push_back((int32_t)gn.size());
      }
      d.ggPtr = upload(gp); d.ggNode = upload(gn);
      std::vector<int32_t> sp(1, 0), st;
      std::vector<int32_t> av((size_t)4*nf, INT32_MIN);
      for (int f = 0; f < nf; ++f) {
        const int L = m.faceL[f], R = m.faceR[f];
        const bool isB = f >= ni;
        const int32_t* fnB = m.faceNodes + m.faceNodePtr[f];
        const int32_t* fnE = m.faceNodes + m.faceNodePtr[f+1];
        const size_t st0 = st.size();
        if (isB) {
          for (int k = m.cellFacePtr[L]; k < m.cellFacePtr[L+1]; ++k) {
            const int cf = m.cellFaces[k];
            if (cf == f) continue;
            bool shares = false;
            for (const int32_t* a = m.faceNodes + m.faceNodePtr[cf]; a != m.faceNodes + m.faceNodePtr[cf+1] && !shares; ++a)
              shares = std::find(fnB, fnE, *a) != fnE;
            if (shares) st.push_back(m.faceL[cf] != L ? m.faceL[cf] : m.faceR[cf]);
          }
          sp.push_back((int32_t)st.size());
        }
        const int32_t* cnL = m.cellNodes + m.cellNodePtr[L];
        const int nL = m.cellNodePtr[L+1] - m.cellNodePtr[L];
        int cnt = 0;
        for (int iNode = 0; iNode < nL; ++iNode) {
          if (std::find(fnB, fnE, cnL[iNode]) == fnE) continue;
          const int j = iNode + 1;                      // the entry after this node in [L nodes | R nodes or stencil]
          int32_t code;
          if (j < nL) code = cnL[j];
          else if (!isB) {
            const int k = j - nL, nR = m.cellNodePtr[R+1] - m.cellNodePtr[R];
            if (k >= nR) throw std::invalid_argument("CorrectedGG: average state reads past the control volume");
            code = m.cellNodes[m.cellNodePtr[R] + k];
          } else {
            const size_t k = (size_t)(j - nL);
            if (st0 + k >= st.size()) throw std::invalid_argument("CorrectedGG: average state reads past the control volume");
            code = ~st[st0 + k];
          }
          if (cnt >= 4) throw std::invalid_argument("CorrectedGG: face with more than 4 nodes");
          av[(size_t)cnt*nf + f] = code; ++cnt;
        }
      }
      d.bndStPtr = upload(sp); d.bndSt = upload(st);
      d.ggAv = upload(av);
    }
    d.saFace = nullptr;
    if (cfg.fluxSplitter == CFB200_FS_ROESA) {
      // RoeSAFlux::updateEtaPA (RoeSAFlux.cxx:146-190): per face and side, the faces of that cell which share a node with
      // the face (every face of the cell with oldStencil); the face itself first.  Partition faces do not exist here: the
      // two overlap layers keep them away from every face that has an updatable side.
      constexpr int W = CFB_SA_WIDTH;
      std::vector<int32_t> sa((size_t)2*W*nf, -1);
      for (int f = 0; f < nf; ++f) {
        const int32_t* fnB = m.faceNodes + m.faceNodePtr[f];
        const int32_t* fnE = m.faceNodes + m.faceNodePtr[f+1];
        for (int side = 0; side < 2; ++side) {
          const int cell = side ? m.faceR[f] : m.faceL[f];
          if (cell >= nc) continue;                         // ghost: RoeSAFlux.cxx:92-94
          int n = 0;
          sa[(size_t)(side*W + n++)*nf + f] = f;
          for (int k = m.cellFacePtr[cell]; k < m.cellFacePtr[cell+1]; ++k) {
            const int g = m.cellFaces[k];
            if (g == f) continue;
            bool take = cfg.roeSaOldStencil != 0;
            for (const int32_t* a = m.faceNodes + m.faceNodePtr[g]; a != m.faceNodes + m.faceNodePtr[g+1] && !take; ++a)
              take = std::find(fnB, fnE, *a) != fnE;
            if (!take) continue;
            if (n >= W) throw std::invalid_argument("RoeSA: a cell with more than 6 faces");
            sa[(size_t)(side*W + n++)*nf + f] = g;
          }
        }
      }
      d.saFace = upload(sa);
    }
    d.bcWallRatio = nullptr;
    {
      bool any = false;
      for (int i = 0; i < nbc; ++i) any = any || (bcs[i].type == CFB200_BC_NOSLIP_ISOTHERMAL);
      if (any) {
        // FVMCC_BaseBC::computeGhostPosition (FVMCC_BaseBC.ci:271-350) per boundary face: the inner node mirrored through
        // the face plane (normal through the face mid point), the wall point half way, then |x_I x_G| / |x_I x_W|
        std::vector<double> ratio(ng, 2.0);
        for (int g = 0; g < ng; ++g) {
          const int f = ni + g;
          const double* n = m.normals + (size_t)f*D;
          const double* xI = m.centres + (size_t)m.faceL[f]*D;
          double mid[3] = {0., 0., 0.};
          const int b = m.faceNodePtr[f], e = m.faceNodePtr[f+1];
          for (int j = b; j < e; ++j) for (int k = 0; k < D; ++k) mid[k] += m.nodes[(size_t)m.faceNodes[j]*D + k];
          for (int k = 0; k < D; ++k) mid[k] /= (double)(e - b);
          double nm = 0., n2 = 0., nx = 0.;
          for (int k = 0; k < D; ++k) { nm += n[k]*mid[k]; n2 += n[k]*n[k]; nx += n[k]*xI[k]; }
          const double t = (nx + (-nm))/n2;
          double dw = 0., dg = 0.;
          for (int k = 0; k < D; ++k) {
            const double xG = xI[k] - 2.*t*n[k], xM = 0.5*(xI[k] + xG);
            dw += (xI[k] - xM)*(xI[k] - xM); dg += (xI[k] - xG)*(xI[k] - xG);
          }
          ratio[g] = std::sqrt(dg)/std::sqrt(dw);
        }
        d.bcWallRatio = upload(ratio);
      }
    }
    d.faceBC = upload(m.faceBC, nf);
    d.updatable = upload(m.updatable, nc);
    for (int f = ni; f < nf; ++f)
      if (m.faceBC[f] < 0 || m.faceBC[f] >= nbc) throw std::invalid_argument("boundary face without a boundary condition");
    d.bcs = upload(bcs, (size_t)nbc);
    hGhostBC.assign(m.faceBC + ni, m.faceBC + nf);      // boundary-condition record of every ghost

    // cell -> faces sorted in loop order, R side flagged by bitwise not (gather order = reference face loop order)
    {
      std::vector<int32_t> sorted(ncf);
      std::vector<uint8_t> active(nc, 0);
      std::vector<int32_t> tmp;
      for (int c = 0; c < nc; ++c) {
        tmp.assign(m.cellFaces + m.cellFacePtr[c], m.cellFaces + m.cellFacePtr[c+1]);
        std::sort(tmp.begin(), tmp.end(), [&](int a, int b) { return fref(a) < fref(b); });     // the caller's loop order
        for (size_t k = 0; k < tmp.size(); ++k) {
          const int f = tmp[k];
          const bool isR = (m.faceL[f] != c);
          sorted[m.cellFacePtr[c] + k] = isR ? ~f : f;
          const bool upL = m.updatable[m.faceL[f]];
          const bool upR = f < ni && m.updatable[m.faceR[f]];
          if (upL || upR) active[c] = 1;
        }
      }
      d.cellFaceSorted = upload(sorted);
      d.cellActive = upload(active);
    }
    // LSQ gather list: replay the reference's edge loop (LeastSquareP1PolyRec2D.cxx:100-134): cell i visits
    // stencil entries that are ghosts or have a higher ID; the edge contributes to i and (if not ghost) to `last`.
    {
      // (cells are visited in the CALLER's order and compared by the caller's IDs: r = reference ID, i = storage ID)
      std::vector<int32_t> cnt(nc + 1, 0);
      for (int r = 0; r < nc; ++r) {
        const int i = cinv(r);
        for (int k = m.stencilPtr[i]; k < m.stencilPtr[i+1]; ++k) {
          const int last = m.stencil[k];
          if (last >= nc || cref(last) > r) { cnt[i+1]++; if (last < nc) cnt[last+1]++; }
        }
      }
      for (int i = 0; i < nc; ++i) cnt[i+1] += cnt[i];
      std::vector<int32_t> pos(cnt.begin(), cnt.end() - 1), nbr(cnt[nc]);
      for (int r = 0; r < nc; ++r) {
        const int i = cinv(r);
        for (int k = m.stencilPtr[i]; k < m.stencilPtr[i+1]; ++k) {
          const int last = m.stencil[k];
          if (last >= nc || cref(last) > r) { nbr[pos[i]++] = last; if (last < nc) nbr[pos[last]++] = i; }
        }
      }
      d.lsqPtr = upload(cnt);
      d.lsqNbr = upload(nbr);
      d.lsqW = dalloc<double>(nbr.size());
    }
    // BSR pattern (FVMCCSparsity.cxx:47-106: diagonal + one block per interior-face neighbour) and face slots
    {
      hRowPtr.assign(1, 0);
      hColInd.clear();
      hColInd.reserve(ncf + nc);
      // rows and columns are the CALLER's cell IDs (the matrix is the consumer's): row r = storage cell cinv(r)
      std::vector<int32_t> cols;
      for (int r = 0; r < nc; ++r) {
        const int c = cinv(r);
        cols.assign(1, r);
        for (int k = m.cellFacePtr[c]; k < m.cellFacePtr[c+1]; ++k) {
          const int f = m.cellFaces[k];
          if (f < ni) cols.push_back(cref(m.faceL[f] == c ? m.faceR[f] : m.faceL[f]));
        }
        std::sort(cols.begin(), cols.end());
        hColInd.insert(hColInd.end(), cols.begin(), cols.end());
        hRowPtr.push_back((int32_t)hColInd.size());
      }
      nnzb = (int64_t)hColInd.size();
      auto slot = [&](int cRow, int cCol) {        // storage cell ids in, slot of block (row, column) out
        const int r = cref(cRow), cidx = cref(cCol);
        const int32_t* b = &hColInd[hRowPtr[r]]; const int32_t* e = &hColInd[hRowPtr[r+1]];
        const int32_t* it = std::lower_bound(b, e, cidx);
        if (it == e || *it != cidx) throw std::runtime_error("BSR slot not found");
        return (int32_t)(it - &hColInd[0]);
      };
      std::vector<int32_t> sLR(ni), sRL(ni), dg(nc);
      for (int f = 0; f < ni; ++f) { sLR[f] = slot(m.faceL[f], m.faceR[f]); sRL[f] = slot(m.faceR[f], m.faceL[f]); }
      for (int c = 0; c < nc; ++c) dg[c] = slot(c, c);
      d.bsrRowPtr = upload(hRowPtr); d.bsrColInd = upload(hColInd);
      d.slotLR = upload(sLR); d.slotRL = upload(sRL); d.diagSlot = upload(dg);
      // diagonal gather sources, in the cell's face-loop order: inner face, cell == L: block(R,cell); cell == R:
      // block(L,cell); boundary face g: ~g
      std::vector<int32_t> dsrc(ncf);
      for (int c = 0; c < nc; ++c) {
        std::vector<int32_t> fs(m.cellFaces + m.cellFacePtr[c], m.cellFaces + m.cellFacePtr[c+1]);
        std::sort(fs.begin(), fs.end(), [&](int a, int b) { return fref(a) < fref(b); });
        for (size_t k = 0; k < fs.size(); ++k) {
          const int f = fs[k];
          dsrc[m.cellFacePtr[c] + k] = (f >= ni) ? ~(f - ni) : ((m.faceL[f] == c) ? sRL[f] : sLR[f]);
        }
      }
      d.diagSrc = upload(dsrc);
      std::vector<int32_t> nu;
      for (int c = 0; c < nc; ++c) if (!m.updatable[c]) nu.push_back(c);
      nNonUpd = (int)nu.size();
      nonUpdCells = upload(nu);
    }
    // fields
    d.faceCentres = dalloc<double>((size_t)nf*D);
    d.lsqL = dalloc<double>((size_t)6*nc);
    d.U = dalloc<double>((size_t)Q*ns);
    d.pastU = dalloc<double>((size_t)Q*nc);
    d.uX = dalloc<double>((size_t)Q*nc); d.uY = dalloc<double>((size_t)Q*nc); d.uZ = dalloc<double>((size_t)Q*nc);
    d.lim = dalloc<double>((size_t)Q*nc);
    d.nstates = dalloc<double>((size_t)Q*nn);
    d.faceFlux = dalloc<double>((size_t)Q*nf);
    d.ucL = dalloc<double>(nf); d.ucR = dalloc<double>(nf); d.ucD = dalloc<double>(nf); d.faceMu = dalloc<double>(nf);
    d.rhs = dalloc<double>((size_t)Q*nc);
    d.updateCoeff = dalloc<double>(nc);
    d.aos = dalloc<double>((size_t)Q*std::max(ns, nn));
    d.partial = dalloc<double>((size_t)nc + (size_t)(nc + 1023)/1024 + 2048);   // + both levels of sumOrdered's block sums
    d.status = dalloc<int>(1);
    d.bndBlocks = dalloc<double>((size_t)ng*Q*Q);
    d.bsrVal = nullptr;       // allocated on the first Jacobian evaluation
    bk.memset(d.uX, 0, sizeof(double)*Q*nc); bk.memset(d.uY, 0, sizeof(double)*Q*nc); bk.memset(d.uZ, 0, sizeof(double)*Q*nc);
    bk.memset(d.updateCoeff, 0, sizeof(double)*nc);
    bk.memset(d.U, 0, sizeof(double)*Q*ns);
    bk.memset(d.nstates, 0, sizeof(double)*Q*nn);
    bk.memset(d.rhs, 0, sizeof(double)*Q*nc);
    bk.memset(d.faceMu, 0, sizeof(double)*nf);
    {
      std::vector<double> ones((size_t)Q*nc, 1.0);   // limiter socket starts at 1 (LeastSquareP1Setup.cxx:127-129)
      bk.h2d(d.lim, ones.data(), ones.size()*sizeof(double));
    }
    d.cellRef = nullptr;
    {
      std::vector<uint8_t> byRow(nc);
      for (int c = 0; c < nc; ++c) byRow[cref(c)] = m.updatable[c];
      dUpdatableByRow = upload(byRow);
    }
    if (cfg.renumber) {
      std::vector<int32_t> sref(ns);
      for (int i = 0; i < ns; ++i) sref[i] = i < nc ? hCellRef[i] : i;
      dStateRef = upload(sref);
      dFaceRef = upload(hFaceRef);
      d.cellRef = dStateRef;
    }
    // split point of the two-part evaluation (computeRhs_ with host copies): the longest prefix of cells whose interior
    // faces all lie in the first half of the interior faces
    splitFace = 0; splitCells = 0;
    if (ni >= 2) {
      const int S = ni/2;
      int n = 0;
      for (; n < nc; ++n) {
        bool ready = true;
        for (int k = m.cellFacePtr[n]; k < m.cellFacePtr[n+1] && ready; ++k) { const int f = m.cellFaces[k]; ready = (f >= ni) || (f < S); }
        if (!ready) break;
      }
      if (n >= nc/4) { splitFace = S; splitCells = n; }
    }
    bk.launch("faceCentres", FaceCentres{d}, nf, 256);
    if (D == 2) bk.launch("lsqSetup", LsqSetup<2>{d}, nc, 256); else bk.launch("lsqSetup", LsqSetup<3>{d}, nc, 256);
    bk.sync();
    meshSet = true;
  }

  void ensureJacobian() {
    if (!d.bsrVal) { d.bsrVal = dalloc<double>((size_t)nnzb*q()*q()); bk.memset(d.bsrVal, 0, sizeof(double)*nnzb*q()*q()); }
  }

  // ---- AoS <-> SoA transfers ---------------------------------------------------------------------
  // ref: storage -> caller index of the items (dStateRef for cell / state arrays, dFaceRef for face arrays) or null
  void putAoS(double* dstSoA, int ld, const double* hostAoS, int n, int k, const int* ref = nullptr) {
    bk.h2d(d.aos, hostAoS, (size_t)n*k*sizeof(double));
    bk.launch("aosToSoa", AosToSoa{d.aos, dstSoA, n, k, ld, ref}, n*k, 256);
    bk.sync();
  }
  void getAoS(const double* srcSoA, int ld, double* hostAoS, int n, int k, const int* ref = nullptr) {
    bk.launch("soaToAos", SoaToAos{srcSoA, d.aos, n, k, ld, ref}, n*k, 256);
    bk.d2h(hostAoS, d.aos, (size_t)n*k*sizeof(double));
  }

  // ---- execute ------------------------------------------------------------------------------------
  template <int D, class PH> void applyBC_() { if (d.nGhosts) bk.launch("applyBC", GhostFill<D, PH>{d, cfg}, d.nGhosts, 128); }

  // the face kernels of the Euler / Navier-Stokes families; DV = derivative computer of the diffusive flux
  // part 0: all faces; part 1: the interior faces [0, splitFace) + the boundary faces; part 2: the interior faces
  // [splitFace, nInner) (the interior kernel is the first launch of every branch: it takes the launch window)
  template <int D, int DV>
  void faceKernels_(bool doJac, int faceShape, int part = 0) {
    const int fam = splitterFamily(cfg);
    if (part == 1) bk.window(0, splitFace); else if (part == 2) bk.window(splitFace, d.nInner);
    if (doJac && cfg.computeJacobian == 2) {
      bk.launchStaged("innerFaceJacAnalytic", AnalyticInnerFaceK<D, DV>{d, cfg}, d.nInner, faceShape);
      if (part != 2) bk.launch("boundaryFaceJac", BoundaryFaceK<D, true, DV>{d, cfg}, d.nGhosts, 128);
    } else if (doJac && (fam == 0 || fam == 3) && roePipeline == 2 && DV == 0) {
      // Roe families, DiamondVolume: two phases, three resident blocks per SM
      const bool cons = (cfg.updateVar == CFB200_UV_CONS);
      if (fam == 0 && cons) bk.launchPhased("innerFaceJac", InnerFacePhased<D, 0, CFB200_UV_CONS>{d, cfg}, d.nInner);
      else if (fam == 0) bk.launchPhased("innerFaceJac", InnerFacePhased<D, 0, CFB200_UV_PVT>{d, cfg}, d.nInner);
      else if (cons) bk.launchPhased("innerFaceJac", InnerFacePhased<D, 3, CFB200_UV_CONS>{d, cfg}, d.nInner);
      else bk.launchPhased("innerFaceJac", InnerFacePhased<D, 3, CFB200_UV_PVT>{d, cfg}, d.nInner);
      if (part != 2) bk.launch("boundaryFaceJac", BoundaryFaceK<D, true, DV>{d, cfg}, d.nGhosts, 128);
    } else if (doJac && (fam == 0 || fam == 3) && roePipeline) {
      // Roe families: update-variable set fixed at compile time, software-pipelined finite-difference loop
      const bool cons = (cfg.updateVar == CFB200_UV_CONS);
      if (fam == 0 && cons) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 0, DV, CFB200_UV_CONS>{d, cfg}, d.nInner, faceShape);
      else if (fam == 0) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 0, DV, CFB200_UV_PVT>{d, cfg}, d.nInner, faceShape);
      else if (cons) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 3, DV, CFB200_UV_CONS>{d, cfg}, d.nInner, faceShape);
      else bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 3, DV, CFB200_UV_PVT>{d, cfg}, d.nInner, faceShape);
      if (part != 2) bk.launch("boundaryFaceJac", BoundaryFaceK<D, true, DV>{d, cfg}, d.nGhosts, 128);
    } else if (doJac) {
      if (fam == 0) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 0, DV>{d, cfg}, d.nInner, faceShape);
      else if (fam == 1) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 1, DV>{d, cfg}, d.nInner, faceShape);
      else if (fam == 2) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 2, DV>{d, cfg}, d.nInner, faceShape);
      else if (fam == 4) bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 4, DV>{d, cfg}, d.nInner, faceShape);
      else bk.launchStaged("innerFaceJac", InnerFaceK<D, true, 3, DV>{d, cfg}, d.nInner, faceShape);
      if (part != 2) bk.launch("boundaryFaceJac", BoundaryFaceK<D, true, DV>{d, cfg}, d.nGhosts, 128);
    } else {
      if (fam == 0) bk.launch("innerFace", InnerFaceK<D, false, 0, DV>{d, cfg}, d.nInner, 128);
      else if (fam == 1) bk.launch("innerFace", InnerFaceK<D, false, 1, DV>{d, cfg}, d.nInner, 128);
      else if (fam == 2) bk.launch("innerFace", InnerFaceK<D, false, 2, DV>{d, cfg}, d.nInner, 128);
      else if (fam == 4) bk.launch("innerFace", InnerFaceK<D, false, 4, DV>{d, cfg}, d.nInner, 128);
      else bk.launch("innerFace", InnerFaceK<D, false, 3, DV>{d, cfg}, d.nInner, 128);
      if (part != 2) bk.launch("boundaryFace", BoundaryFaceK<D, false, DV>{d, cfg}, d.nGhosts, 128);
    }
  }

  // hostRhs / hostUc (optional): host copies of rhs / updateCoeff, started as soon as CellGather has produced them and
  // overlapped with the diagonal gather (the callers of the reference read them right after ComputeRHS: UpdateSol.cxx:83)
  template <int D, class PH> void computeRhs_(bool applyBC, double* hostRhs, double* hostUc) {
    const bool doJac = cfg.computeJacobian != 0;
    if (doJac) ensureJacobian();
    bk.resetTimings();
    bk.memset(d.status, 0, sizeof(int));
    int g0 = 0;
    bool listed = false;
    if (haloPending) {
      // overlap with the halo exchange: ghosts + gradients of the cells that do not see a received state
      if (applyBC) applyBC_<D, PH>();
      if (cfg.polyRec == CFB200_PR_LINEARLS && indepCells && nIndepList > 0) {
        bk.launch("gradientsInterior", Gradients<D, PH>{d, 0, indepCells}, nIndepList, 128);
        listed = true;
      } else if (cfg.polyRec == CFB200_PR_LINEARLS && nIndep > 0) {
        g0 = std::min(nIndep, d.nCells);
        bk.launch("gradientsInterior", Gradients<D, PH>{d, 0, nullptr}, g0, 128);
      }
      joinHalo();                          // the engine's stream now waits for the unpacked overlap states
    }
    if (applyBC) applyBC_<D, PH>();
    // initializeComputationRHS (FVMCC_ComputeRHS.cxx:690-705)
    if (cfg.polyRec == CFB200_PR_LINEARLS) {
      if (listed) bk.launch("gradients", Gradients<D, PH>{d, 0, depCells}, d.nCells - nIndepList, 128);
      else bk.launch("gradients", Gradients<D, PH>{d, g0, nullptr}, d.nCells - g0, 128);
    }
    bk.launch("nodalExtrapolate", NodalExtrapolate<D, PH>{d}, d.nNodes, 256);
    if (cfg.polyRec == CFB200_PR_LINEARLS && cfg.limiter != CFB200_LIM_NONE)
      bk.launch("limiter", LimiterK<D, PH>{d, cfg}, d.nCells, 128);
    const bool fused = (cfg.fuseTimeRhs != 0);
    // With host copies requested the face kernel and the residual gather run in two parts: the cells [0, splitCells) have
    // all their faces among the first splitFace interior faces (+ boundary faces), so their rhs / updateCoeff are final
    // after part 1 and travel to the host while part 2 computes; the rest of the copy runs under the diagonal gather.
    const bool split = splitD2H && splitFace > 0 && (hostRhs || hostUc) && !cfg.renumber;
    auto faces = [&](int part) {
    if constexpr (PH::CENTRED_ONLY) {
      if (part == 1) bk.window(0, splitFace); else if (part == 2) bk.window(splitFace, d.nInner);
      if (doJac) {
        bk.launchStaged2("innerFaceJac", CentredInnerFaceK<PH, true>{d, cfg}, d.nInner);
        if (part != 2) bk.launch("boundaryFaceJac", CentredBoundaryFaceK<PH, true>{d, cfg}, d.nGhosts, 128);
      } else {
        bk.launch("innerFace", CentredInnerFaceK<PH, false>{d, cfg}, d.nInner, 128);
        if (part != 2) bk.launch("boundaryFace", CentredBoundaryFaceK<PH, false>{d, cfg}, d.nGhosts, 128);
      }
    } else {
      if (cfg.derivativeStrategy != CFB200_DV_DIAMOND && cfg.diffusive == CFB200_DIFF_NS) faceKernels_<D, 1>(doJac, faceShape, part);   // Corrected / StateDiff / CorrectedGG: CorrectedWork
      else faceKernels_<D, 0>(doJac, faceShape, part);
    }
    };
    auto gather = [&](int c0, int c1) {      // residual gather + host copies of the cells [c0, c1)
      if (c1 <= c0) return;
      if (c0 > 0 || c1 < d.nCells) bk.window(c0, c1);
      bool srcDone = false;
      if constexpr (PH::Q == 9) {          // Data.SourceTerm (MHD3DProjection only): the variant that adds it at the cell's first face
        if (cfg.sourceTerm != CFB200_ST_NONE) {
          if (fused) bk.launchCellStaged("cellGather", CellGather<D, PH, true, true>{d, cfg}, d.nCells);
          else bk.launch("cellGather", CellGather<D, PH, false, true>{d, cfg}, d.nCells, 256);
          srcDone = true;
        }
      }
      if (srcDone) {}
      else if (fused) bk.launchCellStaged("cellGather", CellGather<D, PH, true>{d, cfg}, d.nCells);
      else bk.launch("cellGather", CellGather<D, PH, false>{d, cfg}, d.nCells, 256);
      if (hostRhs) bk.d2hOverlapped(hostRhs + (size_t)c0*PH::Q, d.rhs + (size_t)c0*PH::Q, sizeof(double)*(size_t)(c1 - c0)*PH::Q);
      if (hostUc) bk.d2hOverlapped(hostUc + c0, d.updateCoeff + c0, sizeof(double)*(size_t)(c1 - c0));
    };
    if (split) {
      faces(1); gather(0, splitCells);
      faces(2); gather(splitCells, d.nCells);
    } else {
      faces(0); gather(0, d.nCells);
    }
    if (doJac) {
      // one thread per (cell, block entry): launched in chunks of < 2^31 entries (q = 9: 26.5 M cells per chunk)
      const int CH = (1 << 30)/DiagGather<D, PH>::GS;          // cells per launch
      for (int c0 = 0; c0 < d.nCells; c0 += CH) {
        const int ne = std::min(CH, d.nCells - c0)*DiagGather<D, PH>::GS;
        bool srcDone = false;
        if constexpr (PH::Q == 9) {
          if (cfg.sourceTerm != CFB200_ST_NONE) {
            if (fused) bk.launch("diagGather", DiagGather<D, PH, true, true>{d, c0, cfg}, ne, 256);
            else bk.launch("diagGather", DiagGather<D, PH, false, true>{d, c0, cfg}, ne, 256);
            srcDone = true;
          }
        }
        if (srcDone) {}
        else if (fused) bk.launch("diagGather", DiagGather<D, PH, true>{d, c0, cfg}, ne, 256);
        else bk.launch("diagGather", DiagGather<D, PH, false>{d, c0, cfg}, ne, 256);
      }
      for (int c0 = 0; c0 < nNonUpd; c0 += CH)
        bk.launch("zeroRows", ZeroRows<D, PH>{d, nonUpdCells, c0}, std::min(CH, nNonUpd - c0)*ZeroRows<D, PH>::GS, 256);
    }
    timeTermApplied = fused;
    int st = 0;
    bk.d2h(&st, d.status, sizeof(int));   // synchronises
    bk.joinCopies();
    lastStatus = st;
  }
  bool isMhd() const { return cfg.physicalModel == CFB200_PM_MHD3D_PROJECTION; }
  void applyBC() {
    if (isMhd()) applyBC_<3, MhdPhys>(); else if (cfg.dim == 2) applyBC_<2, Phys<2> >(); else applyBC_<3, Phys<3> >();
    bk.sync();
  }
  void computeRhs(bool applyBCFirst, double* hostRhs = nullptr, double* hostUc = nullptr) {
    try {
      if (isMhd()) computeRhs_<3, MhdPhys>(applyBCFirst, hostRhs, hostUc);
      else if (cfg.dim == 2) computeRhs_<2, Phys<2> >(applyBCFirst, hostRhs, hostUc);
      else computeRhs_<3, Phys<3> >(applyBCFirst, hostRhs, hostUc);
    } catch (...) { try { bk.joinCopies(); } catch (...) {} throw; }
  }
  template <int D, class PH> void timeRhs_() {
    const int doJac = cfg.computeJacobian && d.bsrVal;
    double minDt = 0.;
    if (cfg.timeUseGlobalDT) {            // FVMCC_PseudoSteadyTimeRhs.cxx:121-127: minimum over every local state
      bk.launch("timeLocalDt", TimeLocalDt<D, PH>{d, cfg}, d.nCells, 256);
      minDt = bk.minOrdered(d.partial, d.nCells);
    }
    bk.launch("timeRhs", TimeRhs<D, PH>{d, cfg, doJac, minDt}, d.nCells, 128);
    bk.sync();
  }
  void timeRhs() {
    if (timeTermApplied) throw std::runtime_error("the pseudo-steady time term was already applied by compute_rhs (fuseTimeRhs)");
    if (isMhd()) timeRhs_<3, MhdPhys>(); else if (cfg.dim == 2) timeRhs_<2, Phys<2> >(); else timeRhs_<3, Phys<3> >();
  }
  void setIndependentCells(const int32_t* cells, int n) {
    std::vector<uint8_t> in(d.nCells, 0);
    for (int i = 0; i < n; ++i) {
      if (cells[i] < 0 || cells[i] >= d.nCells || in[cells[i]]) throw std::invalid_argument("set_independent_cells: bad or repeated cell id");
      in[cells[i]] = 1;
    }
    std::vector<int32_t> a(n), b;
    for (int i = 0; i < n; ++i) a[i] = cinv(cells[i]);
    std::sort(a.begin(), a.end());
    for (int r = 0; r < d.nCells; ++r) if (!in[r]) b.push_back(cinv(r));
    std::sort(b.begin(), b.end());
    indepCells = upload(a); depCells = upload(b); nIndepList = n;
  }
  // position-dependent Dirichlet data of one boundary-condition record: values[iFace*q + v] for the faces of that
  // record in face-loop order; nullptr reverts the record to its constant
  void setBcFaceValues(int bcIndex, const double* values) {
    const int32_t* hostFaceBC = hGhostBC.data();
    const int Q = q(), ng = d.nGhosts;
    if (!bcFaceVal) {
      bcFaceVal = dalloc<double>((size_t)Q*ng); bcFaceFlag = dalloc<uint8_t>(ng);
      bk.memset(bcFaceVal, 0, sizeof(double)*Q*ng); bk.memset(bcFaceFlag, 0, ng);
      hBcFaceVal.assign((size_t)Q*ng, 0.); hBcFaceFlag.assign(ng, 0);
    }
    int k = 0;
    for (int g = 0; g < ng; ++g) {
      if (hostFaceBC[g] != bcIndex) continue;
      hBcFaceFlag[g] = values ? 1 : 0;
      if (values) for (int v = 0; v < Q; ++v) hBcFaceVal[(size_t)v*ng + g] = values[(size_t)k*Q + v];
      ++k;
    }
    bk.h2d(bcFaceVal, hBcFaceVal.data(), sizeof(double)*Q*ng);
    bk.h2d(bcFaceFlag, hBcFaceFlag.data(), ng);
    d.bcFaceVal = bcFaceVal; d.bcFaceFlag = bcFaceFlag;
  }
  // "DistanceBasedGMove": nodeFlag[nNodes] marks the nodes of the TRSs named by `TRSName` (the ones that belong to them,
  // DistanceBasedExtrapolatorGMove.cxx:92-118), idx / values = `ValuesIdx` / `Values`; nodeFlag = nullptr or n = 0 reverts to
  // the plain distance-based extrapolation.  Nodes keep the caller's numbering under cfg.renumber.
  void setNodalPrescribed(const uint8_t* nodeFlag, int n, const int32_t* idx, const double* values) {
    if (!nodeFlag || n == 0) { d.nodePresc = nullptr; d.nNodePresc = 0; return; }
    for (int i = 0; i < n; ++i) if (idx[i] < 0 || idx[i] >= q()) throw std::invalid_argument("prescribed nodal value index out of range");
    if (!nodePresc) nodePresc = dalloc<uint8_t>(d.nNodes);
    if (!nodePrescIdx) { nodePrescIdx = dalloc<int>(CFB200_MAXEQ); nodePrescVal = dalloc<double>(CFB200_MAXEQ); }
    if (n > CFB200_MAXEQ) throw std::invalid_argument("more prescribed nodal values than equations");
    bk.h2d(nodePresc, nodeFlag, d.nNodes);
    bk.h2d(nodePrescIdx, idx, sizeof(int)*n);
    bk.h2d(nodePrescVal, values, sizeof(double)*n);
    d.nodePresc = nodePresc; d.nodePrescIdx = nodePrescIdx; d.nodePrescVal = nodePrescVal; d.nNodePresc = n;
  }
  double updateSolFwdEuler(double cfl, int monitorVar) {
    if (isMhd()) bk.launch("updateSol", UpdateSolFwdEuler<3, MhdPhys>{d, cfl, monitorVar}, d.nCells, 256);
    else if (cfg.dim == 2) bk.launch("updateSol", UpdateSolFwdEuler<2>{d, cfl, monitorVar}, d.nCells, 256);
    else bk.launch("updateSol", UpdateSolFwdEuler<3>{d, cfl, monitorVar}, d.nCells, 256);
    return bk.sumOrdered(d.partial, d.nCells);
  }
  template <int D, class PH> double updateSolNewton_(const double* alpha, int monitorVar) {
    UpdateSolNewton<D, PH> k;
    k.d = d; k.monitorVar = monitorVar;
    for (int v = 0; v < PH::Q; ++v) k.alpha[v] = alpha ? alpha[v] : 1.;
    bk.launch("updateSolNewton", k, d.nCells, 256);
    return bk.sumOrdered(d.partial, d.nCells);
  }
  double updateSolNewton(const double* alpha, int monitorVar) {
    if (isMhd()) return updateSolNewton_<3, MhdPhys>(alpha, monitorVar);
    if (cfg.dim == 2) return updateSolNewton_<2, Phys<2> >(alpha, monitorVar);
    return updateSolNewton_<3, Phys<3> >(alpha, monitorVar);
  }
};

} // namespace cfb
