// cfb200 — host-side construction of the FVMCC mesh data from an element description (what a CFmesh
// file holds): face numbering and L/R orientation, normals, areas, centres, volumes, ghost-state
// positions, LSQ stencil, nodal-extrapolation neighbours/weights, diamond control volumes.
//
// It produces, entry for entry, what the reference's setup phase leaves in its sockets
// (FVMCC_MeshDataBuilder.cxx:91-499, ComputeFaceNormals*P1.cxx, FVMCCGeoDataComputer.ci:61-210,
// ComputeDummyStates.cxx:62-358, ComputeFaceVertexNeighborsPlusGhost.cxx:46-230,
// NodalStatesExtrapolator.ci:134-237,478-591, DistanceBasedExtrapolator.ci:60-125,
// DiamondVolume3DDerivative.cxx:413-563) so that IDs are the reference's IDs, but with a hash-based face
// matcher and threaded per-cell loops instead of the reference's node->face list scans.  tests/ compares it
// bit for bit with the oracle's literal restatement of the reference algorithm.
#include <cstdint>
#include <cmath>
#include <cstring>
#include <vector>
#include <string>
#include <thread>
#include <algorithm>
#include <stdexcept>
#include "../../include/cfb200.h"

namespace {

enum { SH_TRIAG = 0, SH_QUAD = 1, SH_TETRA = 2, SH_HEXA = 3 };

struct ShapeInfo { int nFaces, nFaceNodes; int dof[6][4]; };
// local face -> local nodes (LocalConnectionDataBuilder.cxx:39-58,111-133,301-327,511-551)
const ShapeInfo SHAPES[4] = {
  {3, 2, {{0,1},{1,2},{2,0}}},
  {4, 2, {{0,1},{1,2},{2,3},{3,0}}},
  {4, 3, {{0,1,2},{0,3,1},{1,3,2},{0,2,3}}},
  {6, 4, {{0,3,2,1},{4,5,6,7},{0,1,5,4},{1,2,6,5},{2,3,7,6},{0,4,7,3}}}
};

template <class F> void parallelFor(int n, F f) {
  unsigned nt = std::thread::hardware_concurrency();
  if (nt == 0) nt = 1;
  if (nt > 32) nt = 32;
  if (n < 20000 || nt == 1) { f(0, 0, n); return; }
  std::vector<std::thread> th;
  const int chunk = (n + (int)nt - 1)/(int)nt;
  for (unsigned t = 0; t < nt; ++t) {
    const int b = (int)t*chunk, e = std::min(n, b + chunk);
    if (b >= e) break;
    th.emplace_back([=]() { f((int)t, b, e); });
  }
  for (auto& x : th) x.join();
}

// the flat mesh of the C ABI indexes its CSR tables with 32-bit offsets (cfb200_mesh): a table that does not fit is an error
// of the caller's mesh size, reported (cfb200_mesh_build returns NULL) instead of wrapping around
inline void checkCsrTotal(const std::vector<int32_t>& counts, const char* what) {
  int64_t total = 0;
  for (int32_t c : counts) total += c;
  if (total > (int64_t)INT32_MAX) throw std::length_error(std::string(what) + " table exceeds 2^31 entries");
}

inline void cross(const double* a, const double* b, double* r) {
  r[0] =  a[1]*b[2] - a[2]*b[1];
  r[1] = -a[0]*b[2] + a[2]*b[0];
  r[2] =  a[0]*b[1] - a[1]*b[0];
}
inline double det3(const double* m) {     // CFMat.hh:536-544
  return m[0]*(m[4]*m[8] - m[5]*m[7]) - m[1]*(m[3]*m[8] - m[5]*m[6]) + m[2]*(m[3]*m[7] - m[4]*m[6]);
}
inline double det4(const double* m) {     // CFMat.hh:551-568
  const double a = m[10]*m[15] - m[14]*m[11], b = m[9]*m[15] - m[13]*m[11], c = m[9]*m[14] - m[13]*m[10];
  const double e = m[8]*m[15] - m[12]*m[11], g = m[8]*m[13] - m[12]*m[9], h = m[8]*m[14] - m[12]*m[10];
  return  m[0]*(m[5]*a - m[6]*b + m[7]*c) - m[1]*(m[4]*a - m[6]*e + m[7]*h)
        + m[2]*(m[4]*b - m[5]*e + m[7]*g) - m[3]*(m[4]*c - m[5]*h + m[6]*g);
}
inline double tetraVol(const double* a, const double* b, const double* c, const double* d) {   // VolumeCalculator.hh:141-168
  double m[16];
  const double* p[4] = {a, b, c, d};
  for (int i = 0; i < 4; ++i) { m[i*4] = p[i][0]; m[i*4+1] = p[i][1]; m[i*4+2] = p[i][2]; m[i*4+3] = 1.0; }
  return -(1./6.)*det4(m);
}
inline double pyramVol(const double* c0, const double* c1, const double* c2, const double* c3, const double* c4) {
  const double partial1 = tetraVol(c0, c1, c3, c4);   // VolumeCalculator.cxx:181-205
  const double partial2 = tetraVol(c1, c2, c3, c4);
  return partial1 + partial2;
}

// open-addressing hash of faces keyed by their sorted node tuple
struct FaceTable {
  std::vector<int32_t> slot;      // face id or -1
  std::vector<int32_t> keys;      // 4 sorted nodes per face (-1 padded)
  uint64_t mask;
  explicit FaceTable(size_t expected) {
    size_t cap = 16;
    while (cap < expected*2) cap <<= 1;
    slot.assign(cap, -1);
    mask = cap - 1;
    keys.reserve(expected*4);
  }
  static uint64_t hash(const int32_t* k) {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 4; ++i) { h ^= (uint64_t)(uint32_t)k[i] + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2); h *= 0xBF58476D1CE4E5B9ull; }
    return h ^ (h >> 31);
  }
  // returns the id of the face with this key, inserting it as `newId` when absent (found=false)
  int32_t findOrInsert(const int32_t* k, int32_t newId, bool& found) {
    uint64_t p = hash(k) & mask;
    for (;;) {
      const int32_t s = slot[p];
      if (s < 0) { slot[p] = newId; keys.insert(keys.end(), k, k + 4); found = false; return newId; }
      const int32_t* o = &keys[(size_t)s*4];
      if (o[0] == k[0] && o[1] == k[1] && o[2] == k[2] && o[3] == k[3]) { found = true; return s; }
      p = (p + 1) & mask;
    }
  }
};
inline void sortedKey(const int32_t* nodes, int n, int32_t* k) {
  for (int i = 0; i < 4; ++i) k[i] = i < n ? nodes[i] : -1;
  // tiny insertion sort, descending so that the -1 padding stays at the end
  for (int i = 1; i < 4; ++i) { const int32_t v = k[i]; int j = i - 1; while (j >= 0 && k[j] < v) { k[j+1] = k[j]; --j; } k[j+1] = v; }
}

} // namespace

struct cfb200_meshdata {
  int dim = 0, nCells = 0, nNodes = 0, nFaces = 0, nInner = 0, nGhosts = 0, nTRS = 0;
  std::vector<double> nodes, centres, normals, areas, volumes, nodeW, cvol, faceCentres;
  std::vector<int32_t> faceL, faceR, faceNodePtr, faceNodes, faceGeoID, faceBC, cellFacePtr, cellFaces;
  std::vector<int32_t> stencilPtr, stencil, nodeNbrPtr, nodeNbr, trsPtr, bsrRowPtr, bsrColInd;
  std::vector<int32_t> cellNodePtr, cellNodes;     // the caller's connectivity, kept for CorrectedGG (cfb200_mesh::cellNodes)
  std::vector<uint8_t> cvSign, updatable;
};

namespace {

void build(cfb200_meshdata& M, int dim, int nNodes, const double* nodes, int nCells, const int32_t* cnPtr,
           const int32_t* cn, const int32_t* shape, int nTRS, const int32_t* trsFacePtr, const int32_t* trsFaceCell,
           const int32_t* trsFaceNodePtr, const int32_t* trsFaceNodes) {
  const int D = dim;
  M.dim = dim; M.nCells = nCells; M.nNodes = nNodes; M.nTRS = nTRS;
  M.nodes.assign(nodes, nodes + (size_t)nNodes*D);
  M.updatable.assign(nCells, 1);
  M.cellNodePtr.assign(cnPtr, cnPtr + nCells + 1);
  M.cellNodes.assign(cn, cn + cnPtr[nCells]);
  for (int e = 0; e < nCells; ++e)
    if (shape[e] < 0 || shape[e] > 3 || cnPtr[e+1] - cnPtr[e] != (shape[e] == SH_TRIAG ? 3 : shape[e] == SH_HEXA ? 8 : 4))
      throw std::invalid_argument("unsupported element");

  // ---- geometric faces in creation order (FVMCC_MeshDataBuilder.cxx:129-365)
  M.cellFacePtr.assign(nCells + 1, 0);
  for (int e = 0; e < nCells; ++e) M.cellFacePtr[e+1] = M.cellFacePtr[e] + SHAPES[shape[e]].nFaces;
  const size_t nCF = M.cellFacePtr[nCells];
  std::vector<int32_t> cfGeo(nCF);
  std::vector<int32_t> geoFirstCell, geoSecondCell, geoFirstLF;
  geoFirstCell.reserve(nCF/2 + 1024); geoSecondCell.reserve(nCF/2 + 1024); geoFirstLF.reserve(nCF/2 + 1024);
  {
    FaceTable table(nCF/2 + nCF/8 + 1024);
    int32_t fnodes[4], key[4];
    for (int e = 0; e < nCells; ++e) {
      const ShapeInfo& si = SHAPES[shape[e]];
      for (int lf = 0; lf < si.nFaces; ++lf) {
        for (int i = 0; i < si.nFaceNodes; ++i) fnodes[i] = cn[cnPtr[e] + si.dof[lf][i]];
        sortedKey(fnodes, si.nFaceNodes, key);
        bool found;
        const int32_t id = table.findOrInsert(key, (int32_t)geoFirstCell.size(), found);
        if (!found) { geoFirstCell.push_back(e); geoSecondCell.push_back(-1); geoFirstLF.push_back(lf); }
        else {
          if (geoSecondCell[id] != -1) throw std::runtime_error("face shared by more than two cells");
          geoSecondCell[id] = e;
        }
        cfGeo[M.cellFacePtr[e] + lf] = id;
      }
    }
  }
  const int nGeo = (int)geoFirstCell.size();

  // ---- loop order: inner faces in creation order (createInnerFacesTRS, :369-499), then each boundary TRS
  std::vector<int32_t> geo2loop(nGeo, -1);
  int nInner = 0;
  for (int g = 0; g < nGeo; ++g) if (geoSecondCell[g] >= 0) geo2loop[g] = nInner++;
  const int nB = trsFacePtr[nTRS];
  const int nFaces = nInner + nB;
  M.nInner = nInner; M.nGhosts = nB; M.nFaces = nFaces;
  M.faceL.resize(nFaces); M.faceR.resize(nFaces); M.faceGeoID.resize(nFaces); M.faceBC.assign(nFaces, -1);
  M.faceNodePtr.assign(nFaces + 1, 0);
  std::vector<int32_t> loopLF(nFaces);     // local face index in the L cell
  for (int g = 0; g < nGeo; ++g) {
    const int lf = geo2loop[g];
    if (lf < 0) continue;
    M.faceL[lf] = geoFirstCell[g]; M.faceR[lf] = geoSecondCell[g]; M.faceGeoID[lf] = g; loopLF[lf] = geoFirstLF[g];
  }
  // boundary TRS faces (createBoundaryFacesTRS / createTopologicalRegionSet, :503-834): a TRS face is given by its
  // inner cell and its nodes; it is matched with that cell's boundary face holding the same node set
  M.trsPtr.assign(1, nInner);
  {
    int32_t key[4], key2[4], fnodes[4];
    int ghost = 0;
    for (int t = 0; t < nTRS; ++t) {
      for (int gf = trsFacePtr[t]; gf < trsFacePtr[t+1]; ++gf, ++ghost) {
        const int cell = trsFaceCell[gf];
        const int nbGeoNodes = trsFaceNodePtr[gf+1] - trsFaceNodePtr[gf];
        if (cell < 0 || cell >= nCells || nbGeoNodes > 4) throw std::invalid_argument("bad TRS face");
        sortedKey(&trsFaceNodes[trsFaceNodePtr[gf]], nbGeoNodes, key);
        const ShapeInfo& si = SHAPES[shape[cell]];
        int hit = -1;
        for (int lf = 0; lf < si.nFaces && hit < 0; ++lf) {
          const int g = cfGeo[M.cellFacePtr[cell] + lf];
          if (geoSecondCell[g] >= 0 || si.nFaceNodes != nbGeoNodes) continue;
          for (int i = 0; i < si.nFaceNodes; ++i) fnodes[i] = cn[cnPtr[cell] + si.dof[lf][i]];
          sortedKey(fnodes, si.nFaceNodes, key2);
          if (key[0] == key2[0] && key[1] == key2[1] && key[2] == key2[2] && key[3] == key2[3]) hit = lf;
        }
        if (hit < 0) throw std::runtime_error("boundary face not found");
        const int lfIdx = nInner + ghost;
        const int g = cfGeo[M.cellFacePtr[cell] + hit];
        if (geo2loop[g] >= 0) throw std::runtime_error("boundary face listed twice");
        geo2loop[g] = lfIdx;
        M.faceL[lfIdx] = cell; M.faceR[lfIdx] = nCells + ghost; M.faceGeoID[lfIdx] = g; M.faceBC[lfIdx] = t;
        loopLF[lfIdx] = hit;
      }
      M.trsPtr.push_back(nInner + trsFacePtr[t+1]);
    }
  }
  // face nodes = the L cell's local face nodes, in its local order
  for (int f = 0; f < nFaces; ++f) M.faceNodePtr[f+1] = M.faceNodePtr[f] + SHAPES[shape[M.faceL[f]]].nFaceNodes;
  M.faceNodes.resize(M.faceNodePtr[nFaces]);
  parallelFor(nFaces, [&](int, int b, int e) {
    for (int f = b; f < e; ++f) {
      const int c = M.faceL[f];
      const ShapeInfo& si = SHAPES[shape[c]];
      for (int i = 0; i < si.nFaceNodes; ++i) M.faceNodes[M.faceNodePtr[f] + i] = cn[cnPtr[c] + si.dof[loopLF[f]][i]];
    }
  });
  M.cellFaces.resize(nCF);
  for (size_t i = 0; i < nCF; ++i) {
    M.cellFaces[i] = geo2loop[cfGeo[i]];
    if (M.cellFaces[i] < 0) throw std::runtime_error("face without TRS (partition face) not supported here");
  }

  // ---- cell centres (SetElementStateCoord.cxx:41-57), volumes (LagrangeShapeFunction*P1::computeVolume)
  M.centres.assign((size_t)(nCells + nB)*D, 0.0);
  M.volumes.resize(nCells);
  parallelFor(nCells, [&](int, int b, int e2) {
    for (int e = b; e < e2; ++e) {
      double t[3] = {0, 0, 0};
      const int nn = cnPtr[e+1] - cnPtr[e];
      for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) for (int d = 0; d < D; ++d) t[d] += nodes[(size_t)cn[k]*D + d];
      for (int d = 0; d < D; ++d) M.centres[(size_t)e*D + d] = t[d]/nn;
      const int32_t* c = &cn[cnPtr[e]];
      auto X = [&](int ln) { return &nodes[(size_t)c[ln]*D]; };
      double vol;
      switch (shape[e]) {
        case SH_TRIAG: {
          double m[9];
          for (int i = 0; i < 3; ++i) { m[i*3] = 1.0; m[i*3+1] = X(i)[0]; m[i*3+2] = X(i)[1]; }
          vol = 0.5*det3(m);
          break;
        }
        case SH_QUAD: {
          const double diagonalsProd = (X(2)[0] - X(0)[0])*(X(3)[1] - X(1)[1]) - (X(2)[1] - X(0)[1])*(X(3)[0] - X(1)[0]);
          vol = 0.5*std::abs(diagonalsProd);
          break;
        }
        case SH_TETRA: vol = tetraVol(X(0), X(1), X(2), X(3)); break;
        default: {   // Tonti-Hassan split, VolumeCalculator.cxx:526-567
          double s61[3], s70[3], s63[3], s20[3], s50[3], s64[3], m[9];
          for (int d = 0; d < 3; ++d) {
            s61[d] = X(6)[d] - X(1)[d]; s70[d] = X(7)[d] - X(0)[d]; s63[d] = X(6)[d] - X(3)[d];
            s20[d] = X(2)[d] - X(0)[d]; s50[d] = X(5)[d] - X(0)[d]; s64[d] = X(6)[d] - X(4)[d];
          }
          for (int d = 0; d < 3; ++d) { m[d] = s61[d] + s70[d]; m[3+d] = s63[d]; m[6+d] = s20[d]; }
          vol = det3(m);
          for (int d = 0; d < 3; ++d) { m[d] = s70[d]; m[3+d] = s63[d] + s50[d]; m[6+d] = s64[d]; }
          vol += det3(m);
          for (int d = 0; d < 3; ++d) { m[d] = s61[d]; m[3+d] = s50[d]; m[6+d] = s64[d] + s20[d]; }
          vol += det3(m);
          vol = vol/12.;
        }
      }
      if (vol < 0.0) vol = 0.0;    // FVMCCGeoDataComputer.ci:205-210
      M.volumes[e] = vol;
    }
  });

  // ---- normals from the face's L cell (= the first cell that visits it), areas, face centres
  M.normals.resize((size_t)nFaces*D);
  M.areas.resize(nFaces);
  M.faceCentres.assign((size_t)nFaces*D, 0.0);
  parallelFor(nFaces, [&](int, int b, int e2) {
    for (int f = b; f < e2; ++f) {
      const int e = M.faceL[f], lf = loopLF[f], sh = shape[e];
      const int32_t* c = &cn[cnPtr[e]];
      auto X = [&](int ln) { return &nodes[(size_t)c[ln]*D]; };
      double* nr = &M.normals[(size_t)f*D];
      if (sh == SH_TRIAG || sh == SH_QUAD) {          // ComputeFaceNormalsTriagP1 / QuadP1
        const int nn = (sh == SH_TRIAG) ? 3 : 4;
        const int a = lf, bb = (lf + 1) % nn;
        nr[0] = -X(a)[1] + X(bb)[1];
        nr[1] = -X(bb)[0] + X(a)[0];
      } else if (sh == SH_TETRA) {                    // ComputeFaceNormalsTetraP1
        static const int t[4][4] = {{1,0,2,0},{3,0,1,0},{3,1,2,1},{2,0,3,0}};
        double v1[3], v2[3], v3[3];
        for (int d = 0; d < 3; ++d) { v1[d] = X(t[lf][0])[d] - X(t[lf][1])[d]; v2[d] = X(t[lf][2])[d] - X(t[lf][3])[d]; }
        cross(v1, v2, v3);
        for (int d = 0; d < 3; ++d) nr[d] = -0.5*v3[d];
      } else {                                        // ComputeFaceNormalsHexaP1.cxx:57-153
        static const int h[6][4] = {{2,0,3,1},{6,4,7,5},{5,0,4,1},{6,1,5,2},{6,3,2,7},{7,0,3,4}};
        double v1[3], v2[3], v3[3];
        for (int d = 0; d < 3; ++d) { v1[d] = X(h[lf][0])[d] - X(h[lf][1])[d]; v2[d] = X(h[lf][2])[d] - X(h[lf][3])[d]; }
        cross(v1, v2, v3);
        if (lf == 0) for (int d = 0; d < 3; ++d) nr[d] = -v3[d]*0.5;
        else for (int d = 0; d < 3; ++d) nr[d] = v3[d]*0.5;
      }
      double sq = 0.;
      for (int d = 0; d < D; ++d) sq += nr[d]*nr[d];
      M.areas[f] = std::sqrt(sq);                     // FVMCCGeoDataComputer.ci:161-173
      const int nb = M.faceNodePtr[f+1] - M.faceNodePtr[f];
      const double inv = 1./nb;
      for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k)
        for (int d = 0; d < D; ++d) M.faceCentres[(size_t)f*D + d] += nodes[(size_t)M.faceNodes[k]*D + d]*inv;
    }
  });

  // ---- ghost states: mirror of the inner centre through the face plane (ComputeDummyStates.cxx:62-131,206-358);
  // 4-node boundary faces first get the average of their four triangle normals (rescaled to the face area)
  for (int f = nInner; f < nFaces; ++f) {
    double* normal = &M.normals[(size_t)f*D];
    const int nb = M.faceNodePtr[f+1] - M.faceNodePtr[f];
    const int32_t* fn = &M.faceNodes[M.faceNodePtr[f]];
    if (nb == 4) {
      auto tri = [&](int i0, int i1, int i2, double* tn) {     // NormalsCalculator.hh:156-170
        const double* c0 = &nodes[(size_t)fn[i0]*3]; const double* c1 = &nodes[(size_t)fn[i1]*3]; const double* c2 = &nodes[(size_t)fn[i2]*3];
        double v1[3], v2[3];
        for (int d = 0; d < 3; ++d) { v1[d] = c1[d] - c0[d]; v2[d] = c2[d] - c1[d]; }
        cross(v1, v2, tn);
        for (int d = 0; d < 3; ++d) tn[d] *= 0.5;
        double sq = 0.; for (int d = 0; d < 3; ++d) sq += tn[d]*tn[d];
        const double a = std::sqrt(sq);
        for (int d = 0; d < 3; ++d) tn[d] /= a;
      };
      double t0[3], t1[3], t2[3], t3[3], nn[3];
      tri(0,1,3,t0); tri(1,2,3,t1); tri(0,1,2,t2); tri(0,2,3,t3);
      for (int d = 0; d < 3; ++d) nn[d] = 0.25*(t0[d] + t1[d] + t2[d] + t3[d]);
      double sq = 0.; for (int d = 0; d < 3; ++d) sq += normal[d]*normal[d];
      const double len = std::sqrt(sq);
      for (int d = 0; d < 3; ++d) nn[d] *= len;
      double ip = 0.; for (int d = 0; d < 3; ++d) ip += normal[d]*nn[d];
      if (ip < 0.) for (int d = 0; d < 3; ++d) normal[d] = (-1.)*nn[d];
      else for (int d = 0; d < 3; ++d) normal[d] = nn[d];
    }
    double mid[3] = {0, 0, 0};
    for (int k = 0; k < nb; ++k) for (int d = 0; d < D; ++d) mid[d] += nodes[(size_t)fn[k]*D + d];
    for (int d = 0; d < D; ++d) mid[d] /= (double)nb;
    double ipm = 0.; for (int d = 0; d < D; ++d) ipm += normal[d]*mid[d];
    const double k = -ipm;
    double n2 = 0.; for (int d = 0; d < D; ++d) n2 += normal[d]*normal[d];
    const double* sc = &M.centres[(size_t)M.faceL[f]*D];
    double ips = 0.; for (int d = 0; d < D; ++d) ips += normal[d]*sc[d];
    const double t = (ips + k)/n2;
    const double fac = 2.;
    for (int d = 0; d < D; ++d) M.centres[(size_t)M.faceR[f]*D + d] = sc[d] - fac*t*normal[d];
  }

  // ---- node -> states incidence (cells ascending, then ghosts in loop order), used by the stencil
  std::vector<int32_t> nsPtr(nNodes + 1, 0), nsList;
  {
    for (int e = 0; e < nCells; ++e) for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) nsPtr[cn[k]+1]++;
    for (int f = nInner; f < nFaces; ++f) for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) nsPtr[M.faceNodes[k]+1]++;
    for (int n = 0; n < nNodes; ++n) nsPtr[n+1] += nsPtr[n];
    nsList.resize(nsPtr[nNodes]);
    std::vector<int32_t> pos(nsPtr.begin(), nsPtr.end() - 1);
    for (int e = 0; e < nCells; ++e) for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) nsList[pos[cn[k]]++] = e;
    for (int f = nInner; f < nFaces; ++f) for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) nsList[pos[M.faceNodes[k]]++] = M.faceR[f];
  }
  // ---- LSQ stencil "FaceVertexPlusGhost" (ComputeFaceVertexNeighborsPlusGhost.cxx:46-230): all states sharing a
  // node with the cell.  Canonical order (the reference's is an unstable-sort artefact): cell nodes in local order,
  // per node the cells ascending then the ghosts, first occurrence kept.
  {
    unsigned nt = std::thread::hardware_concurrency(); if (nt == 0) nt = 1; if (nt > 32) nt = 32;
    std::vector<std::vector<int32_t> > lists(nt);
    std::vector<int32_t> counts(nCells);
    std::vector<int> chunkB(nt, 0), chunkE(nt, 0);
    parallelFor(nCells, [&](int t, int b, int e2) {
      chunkB[t] = b; chunkE[t] = e2;
      std::vector<int32_t>& out = lists[t];
      out.reserve((size_t)(e2 - b)*(D == 3 ? 28 : 10));
      for (int e = b; e < e2; ++e) {
        const size_t start = out.size();
        for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) {
          const int node = cn[k];
          for (int j = nsPtr[node]; j < nsPtr[node+1]; ++j) {
            const int32_t s = nsList[j];
            if (s == e) continue;
            bool seen = false;
            for (size_t p = start; p < out.size(); ++p) if (out[p] == s) { seen = true; break; }
            if (!seen) out.push_back(s);
          }
        }
        counts[e] = (int32_t)(out.size() - start);
      }
    });
    M.stencilPtr.assign(nCells + 1, 0);
    checkCsrTotal(counts, "stencil");
    for (int e = 0; e < nCells; ++e) M.stencilPtr[e+1] = M.stencilPtr[e] + counts[e];
    M.stencil.resize(M.stencilPtr[nCells]);
    for (unsigned t = 0; t < nt; ++t)
      if (chunkE[t] > chunkB[t]) std::memcpy(&M.stencil[M.stencilPtr[chunkB[t]]], lists[t].data(), lists[t].size()*sizeof(int32_t));
  }

  // ---- nodal extrapolation neighbours (NodalStatesExtrapolator.ci:134-237,478-591): a boundary node belongs to
  // the first TRS that lists it and takes the ghosts of that TRS's faces plus the cells whose first matching
  // boundary face (>= dim shared nodes) lies in that TRS; an interior node takes every cell around it.
  {
    std::vector<int32_t> trsID(nNodes, -1), bfPtr(nNodes + 1, 0), bfList;
    for (int f = nInner; f < nFaces; ++f)
      for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) {
        const int n = M.faceNodes[k];
        bfPtr[n+1]++;
        if (trsID[n] == -1) trsID[n] = M.faceBC[f];
      }
    for (int n = 0; n < nNodes; ++n) bfPtr[n+1] += bfPtr[n];
    bfList.resize(bfPtr[nNodes]);
    {
      std::vector<int32_t> pos(bfPtr.begin(), bfPtr.end() - 1);
      for (int f = nInner; f < nFaces; ++f)
        for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) bfList[pos[M.faceNodes[k]]++] = f;
    }
    auto cellCounts = [&](int e, int iNode) -> bool {
      const int node = cn[cnPtr[e] + iNode];
      if (bfPtr[node] == bfPtr[node+1]) return true;
      const int nbNodesInCell = cnPtr[e+1] - cnPtr[e];
      for (int i = bfPtr[node]; i < bfPtr[node+1]; ++i) {
        const int f = bfList[i];
        int countNodes = 1;
        for (int j = 0; j < nbNodesInCell; ++j) {
          if (j == iNode) continue;
          const int other = cn[cnPtr[e] + j];
          for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) if (other == M.faceNodes[k]) countNodes++;
        }
        if (countNodes >= D) return M.faceBC[f] == trsID[node];
      }
      return false;
    };
    M.nodeNbrPtr.assign(nNodes + 1, 0);
    for (int f = nInner; f < nFaces; ++f)
      for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) {
        const int n = M.faceNodes[k];
        if (trsID[n] == M.faceBC[f]) M.nodeNbrPtr[n+1]++;
      }
    std::vector<uint8_t> take(cnPtr[nCells]);
    parallelFor(nCells, [&](int, int b, int e2) {
      for (int e = b; e < e2; ++e)
        for (int i = 0; i < cnPtr[e+1] - cnPtr[e]; ++i) take[cnPtr[e] + i] = cellCounts(e, i) ? 1 : 0;
    });
    for (int e = 0; e < nCells; ++e) for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) if (take[k]) M.nodeNbrPtr[cn[k]+1]++;
    for (int n = 0; n < nNodes; ++n) M.nodeNbrPtr[n+1] += M.nodeNbrPtr[n];
    M.nodeNbr.resize(M.nodeNbrPtr[nNodes]);
    std::vector<int32_t> pos(M.nodeNbrPtr.begin(), M.nodeNbrPtr.end() - 1);
    for (int f = nInner; f < nFaces; ++f)
      for (int k = M.faceNodePtr[f]; k < M.faceNodePtr[f+1]; ++k) {
        const int n = M.faceNodes[k];
        if (trsID[n] == M.faceBC[f]) M.nodeNbr[pos[n]++] = M.faceR[f];
      }
    for (int e = 0; e < nCells; ++e) for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) if (take[k]) M.nodeNbr[pos[cn[k]]++] = e;
    // DistanceBasedExtrapolator.ci:60-125
    M.nodeW.resize(M.nodeNbr.size());
    parallelFor(nNodes, [&](int, int b, int e2) {
      for (int n = b; n < e2; ++n) {
        const double* xn = &nodes[(size_t)n*D];
        double sumWeight = 0.0;
        for (int k = M.nodeNbrPtr[n]; k < M.nodeNbrPtr[n+1]; ++k) {
          const double* xs = &M.centres[(size_t)M.nodeNbr[k]*D];
          double dist = 0.; for (int d = 0; d < D; ++d) { const double diff = xs[d] - xn[d]; dist += diff*diff; }
          dist = std::sqrt(dist);
          const double weight = 1./dist;
          M.nodeW[k] = weight;
          sumWeight += weight;
        }
        for (int k = M.nodeNbrPtr[n]; k < M.nodeNbrPtr[n+1]; ++k) M.nodeW[k] /= sumWeight;
      }
    });
  }

  // ---- diamond control volumes (DiamondVolume3DDerivative.cxx:413-563)
  M.cvol.assign(nFaces, 0.0);
  M.cvSign.assign(nFaces, 0);
  if (D == 3) {
    parallelFor(nFaces, [&](int, int b, int e2) {
      for (int f = b; f < e2; ++f) {
        const int nb = M.faceNodePtr[f+1] - M.faceNodePtr[f];
        const int32_t* fn = &M.faceNodes[M.faceNodePtr[f]];
        const double* cL = &M.centres[(size_t)M.faceL[f]*3];
        const double* cR = &M.centres[(size_t)M.faceR[f]*3];
        const double* x0 = &nodes[(size_t)fn[0]*3]; const double* x1 = &nodes[(size_t)fn[1]*3]; const double* x2 = &nodes[(size_t)fn[2]*3];
        double v0, v1;
        if (nb == 3) { v0 = tetraVol(x0, x1, x2, cL); v1 = tetraVol(x0, x1, x2, cR); }
        else { const double* x3 = &nodes[(size_t)fn[3]*3]; v0 = pyramVol(x0, x1, x2, x3, cL); v1 = pyramVol(x0, x1, x2, x3, cR); }
        if (v0 < 0.0) { M.cvSign[f] = 1; v0 = std::abs(v0); }
        if (v1 < 0.0) v1 = std::abs(v1);
        M.cvol[f] = v0 + v1;
      }
    });
  }

  // ---- Jacobian sparsity (FVMCCSparsity.cxx:47-106)
  M.bsrRowPtr.assign(1, 0);
  M.bsrColInd.reserve(nCF + nCells);
  {
    std::vector<int32_t> cols;
    for (int e = 0; e < nCells; ++e) {
      cols.assign(1, e);
      for (int k = M.cellFacePtr[e]; k < M.cellFacePtr[e+1]; ++k) {
        const int f = M.cellFaces[k];
        if (f < nInner) cols.push_back(M.faceL[f] == e ? M.faceR[f] : M.faceL[f]);
      }
      std::sort(cols.begin(), cols.end());
      M.bsrColInd.insert(M.bsrColInd.end(), cols.begin(), cols.end());
      M.bsrRowPtr.push_back((int32_t)M.bsrColInd.size());
    }
  }
}

// ---- the other stencils of `Setup1.stencil` (a3): 1 "FaceVertex" (ComputeFaceVertexNeighbors.cxx:45-129: the cells around
// the cell's nodes, no ghosts), 2 "FaceEdge" (ComputeFaceEdgeNeighbors.cxx:44-150: cells of higher ID sharing >= 2 nodes),
// 3 "Face" (ComputeFaceNeighbors.cxx:46-108: interior-face neighbours of higher ID in cellFaces order).  1 and 2 in
// ascending ID (the reference iterates a set of State pointers).  A cell that shares k nodes with the cell shows up k
// times among the cells gathered around its nodes (the nodes of an element are distinct): after one sort, run lengths
// give the FaceEdge test without comparing node lists.
int rebuildStencil(cfb200_meshdata& M, int kind) {
  if (kind < 1 || kind > 3) return 1;
  const int nCells = M.nCells, nNodes = M.nNodes, nInner = M.nInner;
  const int32_t* cnPtr = M.cellNodePtr.data();
  const int32_t* cn = M.cellNodes.data();
  std::vector<int32_t> ncPtr, ncList;
  if (kind != 3) {
    ncPtr.assign(nNodes + 1, 0);
    for (int e = 0; e < nCells; ++e) for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) ncPtr[cn[k]+1]++;
    for (int n = 0; n < nNodes; ++n) ncPtr[n+1] += ncPtr[n];
    ncList.resize(ncPtr[nNodes]);
    std::vector<int32_t> pos(ncPtr.begin(), ncPtr.end() - 1);
    for (int e = 0; e < nCells; ++e) for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k) ncList[pos[cn[k]]++] = e;
  }
  unsigned nt = std::thread::hardware_concurrency(); if (nt == 0) nt = 1; if (nt > 32) nt = 32;
  std::vector<std::vector<int32_t> > lists(nt);
  std::vector<int32_t> counts(nCells);
  std::vector<int> chunkB(nt, 0), chunkE(nt, 0);
  parallelFor(nCells, [&](int t, int b, int e2) {
    chunkB[t] = b; chunkE[t] = e2;
    std::vector<int32_t>& out = lists[t];
    std::vector<int32_t> around;
    for (int e = b; e < e2; ++e) {
      const size_t start = out.size();
      if (kind == 3) {
        for (int k = M.cellFacePtr[e]; k < M.cellFacePtr[e+1]; ++k) {
          const int f = M.cellFaces[k];
          if (f < nInner && M.faceL[f] == e) out.push_back(M.faceR[f]);     // L < R by construction (a1)
        }
      } else {
        around.clear();
        for (int k = cnPtr[e]; k < cnPtr[e+1]; ++k)
          for (int j = ncPtr[cn[k]]; j < ncPtr[cn[k]+1]; ++j) around.push_back(ncList[j]);
        std::sort(around.begin(), around.end());
        for (size_t i = 0; i < around.size();) {
          size_t j = i;
          while (j < around.size() && around[j] == around[i]) ++j;
          const int32_t s = around[i];
          if (kind == 1 ? (s != e) : (s > e && j - i >= 2)) out.push_back(s);
          i = j;
        }
      }
      counts[e] = (int32_t)(out.size() - start);
    }
  });
  M.stencilPtr.assign(nCells + 1, 0);
  checkCsrTotal(counts, "stencil");
  for (int e = 0; e < nCells; ++e) M.stencilPtr[e+1] = M.stencilPtr[e] + counts[e];
  M.stencil.assign(M.stencilPtr[nCells], 0);
  for (unsigned t = 0; t < nt; ++t)
    if (chunkE[t] > chunkB[t] && !lists[t].empty())
      std::memcpy(&M.stencil[M.stencilPtr[chunkB[t]]], lists[t].data(), lists[t].size()*sizeof(int32_t));
  return 0;
}

template <class T> const void* vptr(const std::vector<T>& v) { return v.data(); }

} // namespace

namespace { thread_local std::string g_meshError; }     // why the last cfb200_mesh_build of this thread returned NULL

extern "C" {

cfb200_meshdata* cfb200_mesh_build(int dim, int nNodes, const double* nodes, int nCells, const int32_t* cellNodePtr,
                                   const int32_t* cellNodes, const int32_t* cellShape, int nTRS, const int32_t* trsFacePtr,
                                   const int32_t* trsFaceCell, const int32_t* trsFaceNodePtr, const int32_t* trsFaceNodes) {
  if (!(dim == 2 || dim == 3) || !nodes || !cellNodePtr || !cellNodes || !cellShape || !trsFacePtr) {
    g_meshError = "cfb200_mesh_build: dim must be 2 or 3 and the element / TRS tables must not be null";
    return nullptr;
  }
  cfb200_meshdata* M = new cfb200_meshdata();
  try {
    build(*M, dim, nNodes, nodes, nCells, cellNodePtr, cellNodes, cellShape, nTRS, trsFacePtr, trsFaceCell,
          trsFaceNodePtr, trsFaceNodes);
  } catch (const std::exception& e) { g_meshError = e.what(); delete M; return nullptr; }
  g_meshError.clear();
  return M;
}

const char* cfb200_mesh_build_error(void) { return g_meshError.c_str(); }

void cfb200_mesh_sizes(const cfb200_meshdata* M, int64_t* out) {
  out[0] = M->nFaces; out[1] = M->nInner; out[2] = M->nGhosts; out[3] = (int64_t)M->faceNodes.size();
  out[4] = (int64_t)M->cellFaces.size(); out[5] = (int64_t)M->stencil.size(); out[6] = (int64_t)M->nodeNbr.size();
  out[7] = (int64_t)M->bsrColInd.size();
}

const void* cfb200_mesh_ptr(const cfb200_meshdata* M, const char* name) {
  const std::string n(name);
#define G(x) if (n == #x) return vptr(M->x);
  G(nodes) G(centres) G(normals) G(areas) G(volumes) G(nodeW) G(cvol) G(faceCentres)
  G(faceL) G(faceR) G(faceNodePtr) G(faceNodes) G(faceGeoID) G(faceBC) G(cellFacePtr) G(cellFaces)
  G(stencilPtr) G(stencil) G(nodeNbrPtr) G(nodeNbr) G(trsPtr) G(cvSign) G(updatable) G(bsrRowPtr) G(bsrColInd)
#undef G
  return nullptr;
}

int cfb200_mesh_get(const cfb200_meshdata* M, const char* name, void* dst) {
  const std::string n(name);
#define G(x) if (n == #x) { if (!M->x.empty()) std::memcpy(dst, M->x.data(), M->x.size()*sizeof(M->x[0])); return 0; }
  G(nodes) G(centres) G(normals) G(areas) G(volumes) G(nodeW) G(cvol) G(faceCentres)
  G(faceL) G(faceR) G(faceNodePtr) G(faceNodes) G(faceGeoID) G(faceBC) G(cellFacePtr) G(cellFaces)
  G(stencilPtr) G(stencil) G(nodeNbrPtr) G(nodeNbr) G(trsPtr) G(cvSign) G(updatable) G(bsrRowPtr) G(bsrColInd)
#undef G
  return 1;
}

void cfb200_mesh_view(const cfb200_meshdata* M, cfb200_mesh* o) {
  o->dim = M->dim; o->nCells = M->nCells; o->nGhosts = M->nGhosts; o->nNodes = M->nNodes; o->nFaces = M->nFaces; o->nInner = M->nInner;
  o->nodes = M->nodes.data(); o->centres = M->centres.data(); o->faceL = M->faceL.data(); o->faceR = M->faceR.data();
  o->faceNodePtr = M->faceNodePtr.data(); o->faceNodes = M->faceNodes.data(); o->normals = M->normals.data();
  o->areas = M->areas.data(); o->volumes = M->volumes.data(); o->cellFacePtr = M->cellFacePtr.data();
  o->cellFaces = M->cellFaces.data(); o->stencilPtr = M->stencilPtr.data(); o->stencil = M->stencil.data();
  o->nodeNbrPtr = M->nodeNbrPtr.data(); o->nodeNbr = M->nodeNbr.data(); o->nodeW = M->nodeW.data();
  o->cvol = M->cvol.data(); o->cvSign = M->cvSign.data(); o->faceBC = M->faceBC.data(); o->updatable = M->updatable.data();
  o->cellNodePtr = M->cellNodePtr.data(); o->cellNodes = M->cellNodes.data();
}

int cfb200_mesh_set_stencil(cfb200_meshdata* M, int kind) {
  if (!M) return CFB200_ERR_INVALID_ARG;
  if (kind == CFB200_STENCIL_FACE_VERTEX_PLUS_GHOST) return CFB200_ERR_INVALID_ARG;   // what cfb200_mesh_build leaves; not rebuilt
  try { return rebuildStencil(*M, kind) ? CFB200_ERR_INVALID_ARG : CFB200_OK; } catch (const std::exception&) { return CFB200_ERR_INVALID_ARG; }
}

void cfb200_mesh_set_updatable(cfb200_meshdata* M, const uint8_t* updatable) {
  std::memcpy(M->updatable.data(), updatable, M->updatable.size());
}

// Diamond control volumes of the faces of a flat mesh (DiamondVolume3DDerivative.cxx:413-563): what a caller that fills
// cfb200_mesh from the reference's sockets cannot read there (they are private to the DerivativeComputer).
void cfb200_mesh_diamond_volumes(const cfb200_mesh* m, double* cvol, uint8_t* cvSign) {
  for (int f = 0; f < m->nFaces; ++f) { cvol[f] = 0.0; cvSign[f] = 0; }
  if (m->dim != 3) return;
  for (int f = 0; f < m->nFaces; ++f) {
    const int nb = m->faceNodePtr[f+1] - m->faceNodePtr[f];
    const int32_t* fn = &m->faceNodes[m->faceNodePtr[f]];
    const double* cL = &m->centres[(size_t)m->faceL[f]*3];
    const double* cR = &m->centres[(size_t)m->faceR[f]*3];
    const double* x0 = &m->nodes[(size_t)fn[0]*3]; const double* x1 = &m->nodes[(size_t)fn[1]*3];
    const double* x2 = &m->nodes[(size_t)fn[2]*3];
    double v0, v1;
    if (nb == 3) { v0 = tetraVol(x0, x1, x2, cL); v1 = tetraVol(x0, x1, x2, cR); }
    else { const double* x3 = &m->nodes[(size_t)fn[3]*3]; v0 = pyramVol(x0, x1, x2, x3, cL); v1 = pyramVol(x0, x1, x2, x3, cR); }
    if (v0 < 0.0) { cvSign[f] = 1; v0 = std::abs(v0); }
    if (v1 < 0.0) v1 = std::abs(v1);
    cvol[f] = v0 + v1;
  }
}

void cfb200_mesh_free(cfb200_meshdata* M) { delete M; }

} // extern "C"
