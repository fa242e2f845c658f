// ORACLE — TEST INFRASTRUCTURE ONLY (see cfo_types.hpp header).
// Literal restatement of the reference's FVMCC mesh-data construction:
// face numbering and L/R orientation, face normals, areas, cell centres and volumes, ghost-state
// positions, LSQ stencil, nodal-extrapolation neighbours and weights, diamond control volumes,
// Jacobian sparsity.  Deliberately the slow algorithm of the reference (node->face lists scanned
// linearly) so that the product's fast host builder can be checked against it bit for bit.
#pragma once
#include "cfo_physics.hpp"
#include <map>
#include <set>

namespace cfo {

enum Shape { SH_TRIAG = 0, SH_QUAD = 1, SH_TETRA = 2, SH_HEXA = 3 };

struct MeshData {
  int dim = 0, nCells = 0, nNodes = 0, nFaces = 0, nInner = 0, nGhosts = 0, nTRS = 0;
  std::vector<double> nodes, centres, normals, areas, volumes, nodeW, cvol, faceCentres;
  std::vector<int32_t> cellNodePtr, cellNodes, cellShape;
  std::vector<int32_t> faceL, faceR, faceNodePtr, faceNodes, faceGeoID, faceBC;
  std::vector<int32_t> cellFacePtr, cellFaces;
  std::vector<int32_t> stencilPtr, stencil, nodeNbrPtr, nodeNbr;
  std::vector<int32_t> trsPtr;      // [nTRS+1] into loop-face index space
  std::vector<uint8_t> cvSign, updatable;
  std::vector<int32_t> bsrRowPtr, bsrColInd;
};

// LocalConnectionDataBuilder.cxx:39-58,111-133,301-327,511-551
static inline int nbFacesOf(int sh) { return sh == SH_TRIAG ? 3 : (sh == SH_QUAD ? 4 : (sh == SH_TETRA ? 4 : 6)); }
static inline int nbFaceNodesOf(int sh) { return sh == SH_TRIAG ? 2 : (sh == SH_QUAD ? 2 : (sh == SH_TETRA ? 3 : 4)); }
static inline const int* faceDof(int sh, int iFace) {
  static const int tri[3][2] = {{0,1},{1,2},{2,0}};
  static const int quad[4][2] = {{0,1},{1,2},{2,3},{3,0}};
  static const int tet[4][3] = {{0,1,2},{0,3,1},{1,3,2},{0,2,3}};
  static const int hex[6][4] = {{0,3,2,1},{4,5,6,7},{0,1,5,4},{1,2,6,5},{2,3,7,6},{0,4,7,3}};
  switch (sh) { case SH_TRIAG: return tri[iFace]; case SH_QUAD: return quad[iFace];
                case SH_TETRA: return tet[iFace]; default: return hex[iFace]; }
}

static inline double det3(const double* m) {   // CFMat.hh:536-544
  return m[0]*(m[4]*m[8] - m[5]*m[7]) - m[1]*(m[3]*m[8] - m[5]*m[6]) + m[2]*(m[3]*m[7] - m[4]*m[6]);
}
static inline double det4(const double* m) {   // CFMat.hh:551-568
  const double d00 = m[0], d01 = m[1], d02 = m[2], d03 = m[3];
  const double d04 = m[4], d05 = m[5], d06 = m[6], d07 = m[7];
  const double d10d15_d14d11 = m[10]*m[15] - m[14]*m[11];
  const double d09d15_d13d11 = m[ 9]*m[15] - m[13]*m[11];
  const double d09d14_d13d10 = m[ 9]*m[14] - m[13]*m[10];
  const double d08d15_d12d11 = m[ 8]*m[15] - m[12]*m[11];
  const double d08d13_d12d09 = m[ 8]*m[13] - m[12]*m[ 9];
  const double d08d14_d12d10 = m[ 8]*m[14] - m[12]*m[10];
  return  d00*(d05*(d10d15_d14d11) - d06*(d09d15_d13d11) + d07*(d09d14_d13d10))
    - d01*(d04*(d10d15_d14d11) - d06*(d08d15_d12d11) + d07*(d08d14_d12d10))
    + d02*(d04*(d09d15_d13d11) - d05*(d08d15_d12d11) + d07*(d08d13_d12d09))
    - d03*(d04*(d09d14_d13d10) - d05*(d08d14_d12d10) + d06*(d08d13_d12d09));
}
// VolumeCalculator.hh:141-168: -(1/6) det[[xi 1]..]
static inline double tetraVol(const double* a, const double* b, const double* c, const double* d) {
  double m[16];
  const double* p[4] = {a, b, c, d};
  for (int i = 0; i < 4; ++i) { m[i*4+0] = p[i][0]; m[i*4+1] = p[i][1]; m[i*4+2] = p[i][2]; m[i*4+3] = 1.0; }
  return -(1./6.)*det4(m);
}
// VolumeCalculator.cxx:181-205 (RealMatrix overload)
static inline double pyramVol(const double* c0, const double* c1, const double* c2, const double* c3, const double* c4) {
  const double partial1 = tetraVol(c0, c1, c3, c4);
  const double partial2 = tetraVol(c1, c2, c3, c4);
  return partial1 + partial2;
}

inline void buildMesh(int dim, int nNodes, const double* nodes, int nCells, const int32_t* cellNodePtr,
                      const int32_t* cellNodes, const int32_t* cellShape,
                      int nTRS, const int32_t* trsFacePtr /*[nTRS+1]*/, const int32_t* trsFaceCell,
                      const int32_t* trsFaceNodePtr, const int32_t* trsFaceNodes, MeshData& M) {
  M.dim = dim; M.nCells = nCells; M.nNodes = nNodes; M.nTRS = nTRS;
  M.nodes.assign(nodes, nodes + (size_t)nNodes*dim);
  M.cellNodePtr.assign(cellNodePtr, cellNodePtr + nCells + 1);
  M.cellNodes.assign(cellNodes, cellNodes + cellNodePtr[nCells]);
  M.cellShape.assign(cellShape, cellShape + nCells);
  M.updatable.assign(nCells, 1);
  const int D = dim;

  // ---- createCellFaces (FVMCC_MeshDataBuilder.cxx:129-365)
  std::vector<std::vector<int> > mapNodeFace(nNodes);
  std::vector<int> cfPtr(nCells + 1, 0);
  for (int e = 0; e < nCells; ++e) cfPtr[e+1] = cfPtr[e] + nbFacesOf(cellShape[e]);
  std::vector<int> cellFacesGeo(cfPtr[nCells]);
  std::vector<uint8_t> isBFace; isBFace.reserve(cfPtr[nCells]);
  int nbGeoFaces = 0;
  std::vector<int> nodesInFace(8);
  for (int e = 0; e < nCells; ++e) {
    const int sh = cellShape[e];
    const int nbElemFaces = nbFacesOf(sh);
    const int nbNodesPerFace = nbFaceNodesOf(sh);
    for (int iFace = 0; iFace < nbElemFaces; ++iFace) {
      const int* dof = faceDof(sh, iFace);
      for (int i = 0; i < nbNodesPerFace; ++i) nodesInFace[i] = cellNodes[cellNodePtr[e] + dof[i]];
      const int nodeID = nodesInFace[0];
      const int nbFaceIDsRefNode = (int)mapNodeFace[nodeID].size();
      bool faceFound = false;
      for (int iFaceID = 0; iFaceID < nbFaceIDsRefNode && !faceFound; ++iFaceID) {
        const int currFaceID = mapNodeFace[nodeID][iFaceID];
        int countNodes = 1;
        for (int iNode = 1; iNode < nbNodesPerFace; ++iNode) {
          const int newNodeID = nodesInFace[iNode];
          for (int jf : mapNodeFace[newNodeID]) if (jf == currFaceID) { ++countNodes; break; }
          if (countNodes == nbNodesPerFace) {
            faceFound = true;
            cellFacesGeo[cfPtr[e] + iFace] = currFaceID;
            isBFace[currFaceID] = 0;
            break;
          }
        }
      }
      if (!faceFound || nbFaceIDsRefNode == 0) {
        for (int i = 0; i < nbNodesPerFace; ++i) mapNodeFace[nodesInFace[i]].push_back(nbGeoFaces);
        cellFacesGeo[cfPtr[e] + iFace] = nbGeoFaces;
        isBFace.push_back(1);
        nbGeoFaces++;
      }
    }
  }

  // ---- createInnerFacesTRS (FVMCC_MeshDataBuilder.cxx:369-499)
  std::vector<int> idxState(nbGeoFaces, -1);
  std::vector<int> inL, inR, inGeo, inNodePtr(1, 0), inNodes;
  std::vector<int> bCell, bGeo, bNodePtr(1, 0), bNodes;
  for (int e = 0; e < nCells; ++e) {
    const int sh = cellShape[e];
    for (int iFace = 0; iFace < nbFacesOf(sh); ++iFace) {
      const int nbNodesPerFace = nbFaceNodesOf(sh);
      const int* dof = faceDof(sh, iFace);
      const int faceID = cellFacesGeo[cfPtr[e] + iFace];
      if (!isBFace[faceID]) {
        if (idxState[faceID] == -1) {
          for (int i = 0; i < nbNodesPerFace; ++i) inNodes.push_back(cellNodes[cellNodePtr[e] + dof[i]]);
          inNodePtr.push_back((int)inNodes.size());
          inGeo.push_back(faceID); inL.push_back(e); inR.push_back(-1);
          idxState[faceID] = (int)inL.size() - 1;
        } else {
          inR[idxState[faceID]] = e;
        }
      } else {
        for (int i = 0; i < nbNodesPerFace; ++i) bNodes.push_back(cellNodes[cellNodePtr[e] + dof[i]]);
        bNodePtr.push_back((int)bNodes.size());
        bGeo.push_back(faceID); bCell.push_back(e);
      }
    }
  }
  const int nInner = (int)inL.size();
  M.nInner = nInner;

  // ---- createBoundaryFacesTRS / createTopologicalRegionSet (FVMCC_MeshDataBuilder.cxx:503-834)
  const int nB = trsFacePtr[nTRS];
  M.nGhosts = nB;
  M.nFaces = nInner + nB;
  M.faceL = inL; M.faceR = inR; M.faceGeoID = inGeo; M.faceNodePtr = inNodePtr; M.faceNodes = inNodes;
  M.faceBC.assign(nInner, -1);
  M.trsPtr.assign(1, nInner);
  std::vector<std::vector<int> > cellToB(nCells);
  for (int i = 0; i < (int)bCell.size(); ++i) cellToB[bCell[i]].push_back(i);
  int ghost = 0;
  for (int iTRS = 0; iTRS < nTRS; ++iTRS) {
    for (int g = trsFacePtr[iTRS]; g < trsFacePtr[iTRS+1]; ++g, ++ghost) {
      const int stateID = trsFaceCell[g];
      const int nbGeoNodes = trsFaceNodePtr[g+1] - trsFaceNodePtr[g];
      int faceIdx = -1;
      for (int cand : cellToB[stateID]) {
        int nbMatchingNodes = 0;
        if (nbGeoNodes == bNodePtr[cand+1] - bNodePtr[cand]) {
          for (int iNode = 0; iNode < nbGeoNodes; ++iNode) {
            const int nodeID = bNodes[bNodePtr[cand] + iNode];
            for (int jNode = 0; jNode < nbGeoNodes; ++jNode)
              if (trsFaceNodes[trsFaceNodePtr[g] + jNode] == nodeID) { ++nbMatchingNodes; break; }
          }
        }
        if (nbMatchingNodes == nbGeoNodes) { faceIdx = cand; break; }
      }
      if (faceIdx < 0) throw std::runtime_error("boundary face not found");
      for (int iNode = 0; iNode < nbGeoNodes; ++iNode) M.faceNodes.push_back(bNodes[bNodePtr[faceIdx] + iNode]);
      M.faceNodePtr.push_back((int)M.faceNodes.size());
      M.faceL.push_back(stateID); M.faceR.push_back(nCells + ghost);
      M.faceGeoID.push_back(bGeo[faceIdx]); M.faceBC.push_back(iTRS);
    }
    M.trsPtr.push_back((int)M.faceL.size());
  }
  // cellFaces in loop-face index space
  std::vector<int> geo2loop(nbGeoFaces, -1);
  for (int lf = 0; lf < M.nFaces; ++lf) geo2loop[M.faceGeoID[lf]] = lf;
  M.cellFacePtr.assign(cfPtr.begin(), cfPtr.end());
  M.cellFaces.resize(cfPtr[nCells]);
  for (size_t i = 0; i < cellFacesGeo.size(); ++i) {
    M.cellFaces[i] = geo2loop[cellFacesGeo[i]];
    if (M.cellFaces[i] < 0) throw std::runtime_error("face without TRS (partition face) not supported here");
  }

  // ---- cell centres (SetElementStateCoord.cxx:41-57)
  M.centres.assign((size_t)(nCells + nB)*D, 0.0);
  for (int e = 0; e < nCells; ++e) {
    double t[3] = {0,0,0};
    const int nn = cellNodePtr[e+1] - cellNodePtr[e];
    for (int k = cellNodePtr[e]; k < cellNodePtr[e+1]; ++k)
      for (int d = 0; d < D; ++d) t[d] += nodes[(size_t)cellNodes[k]*D + d];
    for (int d = 0; d < D; ++d) M.centres[(size_t)e*D + d] = t[d]/nn;
  }

  // ---- normals (ComputeFaceNormals{Triag,Quad,Tetra,Hexa}P1.cxx) in geo-face space, first visiting cell
  std::vector<double> gnormals((size_t)nbGeoFaces*D, 0.0);
  std::vector<int> isOutward(nbGeoFaces, -1);
  for (int e = 0; e < nCells; ++e) {
    const int sh = cellShape[e];
    const int32_t* cn = &cellNodes[cellNodePtr[e]];
    auto X = [&](int ln) { return &nodes[(size_t)cn[ln]*D]; };
    for (int iFace = 0; iFace < nbFacesOf(sh); ++iFace) {
      const int faceID = cellFacesGeo[cfPtr[e] + iFace];
      if (isOutward[faceID] != -1) continue;
      double* nr = &gnormals[(size_t)faceID*D];
      if (sh == SH_TRIAG || sh == SH_QUAD) {
        const int nn = (sh == SH_TRIAG) ? 3 : 4;
        const int a = iFace, b = (iFace + 1) % nn;
        nr[0] = -X(a)[1] + X(b)[1];
        nr[1] = -X(b)[0] + X(a)[0];
      } else if (sh == SH_TETRA) {
        static const int t[4][4] = {{1,0,2,0},{3,0,1,0},{3,1,2,1},{2,0,3,0}};
        double v1[3], v2[3], v3[3];
        for (int d = 0; d < 3; ++d) { v1[d] = X(t[iFace][0])[d] - X(t[iFace][1])[d]; v2[d] = X(t[iFace][2])[d] - X(t[iFace][3])[d]; }
        crossProd(v1, v2, v3);
        for (int d = 0; d < 3; ++d) nr[d] = -0.5*v3[d];
      } else {
        // ComputeFaceNormalsHexaP1.cxx:57-153
        static const int h[6][4] = {{2,0,3,1},{6,4,7,5},{5,0,4,1},{6,1,5,2},{6,3,2,7},{7,0,3,4}};
        double v1[3], v2[3], v3[3];
        for (int d = 0; d < 3; ++d) { v1[d] = X(h[iFace][0])[d] - X(h[iFace][1])[d]; v2[d] = X(h[iFace][2])[d] - X(h[iFace][3])[d]; }
        crossProd(v1, v2, v3);
        if (iFace == 0) for (int d = 0; d < 3; ++d) nr[d] = -v3[d]*0.5;
        else for (int d = 0; d < 3; ++d) nr[d] = v3[d]*0.5;
      }
      isOutward[faceID] = e;
    }
  }
  // face areas (FVMCCGeoDataComputer.ci:161-173): computed BEFORE the ghost-state pass overrides
  // the normals of 4-node boundary faces
  M.normals.resize((size_t)M.nFaces*D);
  M.areas.resize(M.nFaces);
  for (int lf = 0; lf < M.nFaces; ++lf) {
    double sq = 0.;
    for (int d = 0; d < D; ++d) { const double v = gnormals[(size_t)M.faceGeoID[lf]*D + d]; M.normals[(size_t)lf*D + d] = v; sq += v*v; }
    M.areas[lf] = std::sqrt(sq);
  }

  // ---- ghost states (ComputeDummyStates.cxx:62-131,206-358)
  for (int lf = nInner; lf < M.nFaces; ++lf) {
    double* normal = &M.normals[(size_t)lf*D];
    const int nb = M.faceNodePtr[lf+1] - M.faceNodePtr[lf];
    const int32_t* fn = &M.faceNodes[M.faceNodePtr[lf]];
    if (nb == 4) {
      auto tri = [&](int i0, int i1, int i2, double* tn) {   // NormalsCalculator.hh:156-170
        const double* c0 = &nodes[(size_t)fn[i0]*3]; const double* c1 = &nodes[(size_t)fn[i1]*3]; const double* c2 = &nodes[(size_t)fn[i2]*3];
        double v1[3], v2[3];
        for (int d = 0; d < 3; ++d) { v1[d] = c1[d] - c0[d]; v2[d] = c2[d] - c1[d]; }
        crossProd(v1, v2, tn);
        for (int d = 0; d < 3; ++d) tn[d] *= 0.5;
        double sq = 0.; for (int d = 0; d < 3; ++d) sq += tn[d]*tn[d];
        const double a = std::sqrt(sq);
        for (int d = 0; d < 3; ++d) tn[d] /= a;
      };
      double t0[3], t1[3], t2[3], t3[3], newNormal[3];
      tri(0,1,3,t0); tri(1,2,3,t1); tri(0,1,2,t2); tri(0,2,3,t3);
      for (int d = 0; d < 3; ++d) newNormal[d] = 0.25*(t0[d] + t1[d] + t2[d] + t3[d]);
      double sq = 0.; for (int d = 0; d < 3; ++d) sq += normal[d]*normal[d];
      const double nn2 = std::sqrt(sq);
      for (int d = 0; d < 3; ++d) newNormal[d] *= nn2;
      double ip = 0.; for (int d = 0; d < 3; ++d) ip += normal[d]*newNormal[d];
      if (ip < 0.) for (int d = 0; d < 3; ++d) normal[d] = (-1.)*newNormal[d];
      else for (int d = 0; d < 3; ++d) normal[d] = newNormal[d];
    }
    double mid[3] = {0,0,0};
    for (int k = 0; k < nb; ++k) for (int d = 0; d < D; ++d) mid[d] += nodes[(size_t)fn[k]*D + d];
    for (int d = 0; d < D; ++d) mid[d] /= (double)nb;
    double ipm = 0.; for (int d = 0; d < D; ++d) ipm += normal[d]*mid[d];
    const double k = -ipm;
    double n2 = 0.; for (int d = 0; d < D; ++d) n2 += normal[d]*normal[d];
    const double* sc = &M.centres[(size_t)M.faceL[lf]*D];
    double ips = 0.; for (int d = 0; d < D; ++d) ips += normal[d]*sc[d];
    const double t = (ips + k)/n2;
    const double fac = 2.;
    for (int d = 0; d < D; ++d) M.centres[(size_t)M.faceR[lf]*D + d] = sc[d] - fac*t*normal[d];
  }

  // ---- volumes (LagrangeShapeFunction*P1::computeVolume)
  M.volumes.resize(nCells);
  for (int e = 0; e < nCells; ++e) {
    const int sh = cellShape[e];
    const int32_t* cn = &cellNodes[cellNodePtr[e]];
    auto X = [&](int ln) { return &nodes[(size_t)cn[ln]*D]; };
    double vol;
    if (sh == SH_TRIAG) {
      double m[9];
      for (int i = 0; i < 3; ++i) { m[i*3] = 1.0; m[i*3+1] = X(i)[0]; m[i*3+2] = X(i)[1]; }
      vol = 0.5*det3(m);
    } else if (sh == SH_QUAD) {
      const double diagonalsProd = (X(2)[0] - X(0)[0])*(X(3)[1] - X(1)[1]) - (X(2)[1] - X(0)[1])*(X(3)[0] - X(1)[0]);
      vol = 0.5*std::abs(diagonalsProd);
    } else if (sh == SH_TETRA) {
      vol = tetraVol(X(0), X(1), X(2), X(3));
    } else {   // VolumeCalculator.cxx:526-567 (Tonti-Hassan)
      double s61[3], s70[3], s63[3], s20[3], s50[3], s64[3], tmp[3], det[9];
      for (int d = 0; d < 3; ++d) {
        s61[d] = X(6)[d] - X(1)[d]; s70[d] = X(7)[d] - X(0)[d]; s63[d] = X(6)[d] - X(3)[d];
        s20[d] = X(2)[d] - X(0)[d]; s50[d] = X(5)[d] - X(0)[d]; s64[d] = X(6)[d] - X(4)[d];
      }
      for (int d = 0; d < 3; ++d) { tmp[d] = s61[d] + s70[d]; det[d] = tmp[d]; det[3+d] = s63[d]; det[6+d] = s20[d]; }
      vol = det3(det);
      for (int d = 0; d < 3; ++d) { tmp[d] = s63[d] + s50[d]; det[d] = s70[d]; det[3+d] = tmp[d]; det[6+d] = s64[d]; }
      vol += det3(det);
      for (int d = 0; d < 3; ++d) { tmp[d] = s64[d] + s20[d]; det[d] = s61[d]; det[3+d] = s50[d]; det[6+d] = tmp[d]; }
      vol += det3(det);
      vol = vol/12.;
    }
    if (vol < 0.0) vol = 0.0;   // FVMCCGeoDataComputer.ci:205-210
    M.volumes[e] = vol;
  }

  // ---- face centres (FVMCCGeoDataComputer.ci computeFaceCenters)
  M.faceCentres.assign((size_t)M.nFaces*D, 0.0);
  for (int lf = 0; lf < M.nFaces; ++lf) {
    const int nb = M.faceNodePtr[lf+1] - M.faceNodePtr[lf];
    const double inv = 1./nb;
    for (int k = M.faceNodePtr[lf]; k < M.faceNodePtr[lf+1]; ++k)
      for (int d = 0; d < D; ++d) M.faceCentres[(size_t)lf*D + d] += nodes[(size_t)M.faceNodes[k]*D + d]*inv;
  }

  // ---- stencil (ComputeFaceVertexNeighborsPlusGhost.cxx:46-230).  CANONICAL ORDER: the reference sorts
  // a multimap by node key with an unstable sort, so the order among states sharing a node is
  // implementation-defined; here: cells in ascending ID, then ghosts in TRS order (= insertion order).
  std::vector<std::vector<int> > nodeStates(nNodes);
  for (int e = 0; e < nCells; ++e)
    for (int k = cellNodePtr[e]; k < cellNodePtr[e+1]; ++k) nodeStates[cellNodes[k]].push_back(e);
  for (int lf = nInner; lf < M.nFaces; ++lf)
    for (int k = M.faceNodePtr[lf]; k < M.faceNodePtr[lf+1]; ++k) nodeStates[M.faceNodes[k]].push_back(M.faceR[lf]);
  M.stencilPtr.assign(1, 0);
  std::vector<int> neighborList;
  for (int e = 0; e < nCells; ++e) {
    neighborList.clear();
    for (int k = cellNodePtr[e]; k < cellNodePtr[e+1]; ++k) {
      for (int s : nodeStates[cellNodes[k]]) {
        if (s != e && std::find(neighborList.begin(), neighborList.end(), s) == neighborList.end()) neighborList.push_back(s);
      }
    }
    M.stencil.insert(M.stencil.end(), neighborList.begin(), neighborList.end());
    M.stencilPtr.push_back((int)M.stencil.size());
  }

  // ---- nodal extrapolation neighbours (NodalStatesExtrapolator.ci:134-237,478-591) and weights
  // (DistanceBasedExtrapolator.ci:60-125)
  {
    std::vector<int> trsID(nNodes, -1);
    std::vector<std::vector<int> > bFacesPerNode(nNodes);
    std::vector<std::vector<int> > nb(nNodes);
    for (int iTRS = 0; iTRS < nTRS; ++iTRS) {
      for (int lf = M.trsPtr[iTRS]; lf < M.trsPtr[iTRS+1]; ++lf) {
        for (int k = M.faceNodePtr[lf]; k < M.faceNodePtr[lf+1]; ++k) {
          const int nodeID = M.faceNodes[k];
          bFacesPerNode[nodeID].push_back(lf);
          if (trsID[nodeID] == -1) trsID[nodeID] = iTRS;
          if (trsID[nodeID] == iTRS) nb[nodeID].push_back(M.faceR[lf]);
        }
      }
    }
    for (int e = 0; e < nCells; ++e) {
      const int nbNodesInCell = cellNodePtr[e+1] - cellNodePtr[e];
      for (int iNode = 0; iNode < nbNodesInCell; ++iNode) {
        const int nodeID = cellNodes[cellNodePtr[e] + iNode];
        const int nbBFacesPerNode = (int)bFacesPerNode[nodeID].size();
        if (nbBFacesPerNode > 0) {
          for (int i = 0; i < nbBFacesPerNode; ++i) {
            int countNodes = 1;
            const int lf = bFacesPerNode[nodeID][i];
            for (int jNode = 0; jNode < nbNodesInCell; ++jNode) {
              if (jNode != iNode) {
                const int nodeIDToMatch = cellNodes[cellNodePtr[e] + jNode];
                for (int k = M.faceNodePtr[lf]; k < M.faceNodePtr[lf+1]; ++k) if (nodeIDToMatch == M.faceNodes[k]) countNodes++;
              }
            }
            if (countNodes >= D) {
              if (M.faceBC[lf] == trsID[nodeID]) nb[nodeID].push_back(e);
              break;
            }
          }
        } else {
          nb[nodeID].push_back(e);
        }
      }
    }
    M.nodeNbrPtr.assign(1, 0);
    for (int n = 0; n < nNodes; ++n) {
      const double* xn = &nodes[(size_t)n*D];
      double sumWeight = 0.0;
      const size_t start = M.nodeW.size();
      for (int s : nb[n]) {
        const double* xs = &M.centres[(size_t)s*D];
        double dist = 0.; for (int d = 0; d < D; ++d) { const double diff = xs[d] - xn[d]; dist += diff*diff; }
        dist = std::sqrt(dist);
        const double weight = 1./dist;
        M.nodeNbr.push_back(s); M.nodeW.push_back(weight);
        sumWeight += weight;
      }
      for (size_t k = start; k < M.nodeW.size(); ++k) M.nodeW[k] /= sumWeight;
      M.nodeNbrPtr.push_back((int)M.nodeNbr.size());
    }
  }

  // ---- diamond control volumes (DiamondVolume3DDerivative.cxx:413-563; 2D ones are computed on the fly)
  M.cvol.assign(M.nFaces, 0.0);
  M.cvSign.assign(M.nFaces, 0);
  if (D == 3) {
    for (int lf = 0; lf < M.nFaces; ++lf) {
      const int nb = M.faceNodePtr[lf+1] - M.faceNodePtr[lf];
      const int32_t* fn = &M.faceNodes[M.faceNodePtr[lf]];
      const double* cL = &M.centres[(size_t)M.faceL[lf]*3];
      const double* cR = &M.centres[(size_t)M.faceR[lf]*3];
      if (nb == 3) {
        double v0 = tetraVol(&nodes[(size_t)fn[0]*3], &nodes[(size_t)fn[1]*3], &nodes[(size_t)fn[2]*3], cL);
        if (v0 < 0.0) { M.cvSign[lf] = 1; v0 = std::abs(v0); }
        double v1 = tetraVol(&nodes[(size_t)fn[0]*3], &nodes[(size_t)fn[1]*3], &nodes[(size_t)fn[2]*3], cR);
        if (v1 < 0.0) v1 = std::abs(v1);
        M.cvol[lf] = v0 + v1;
      } else {
        double v0 = pyramVol(&nodes[(size_t)fn[0]*3], &nodes[(size_t)fn[1]*3], &nodes[(size_t)fn[2]*3], &nodes[(size_t)fn[3]*3], cL);
        if (v0 < 0.0) { M.cvSign[lf] = 1; v0 = std::abs(v0); }
        double v1 = pyramVol(&nodes[(size_t)fn[0]*3], &nodes[(size_t)fn[1]*3], &nodes[(size_t)fn[2]*3], &nodes[(size_t)fn[3]*3], cR);
        if (v1 < 0.0) v1 = std::abs(v1);
        M.cvol[lf] = v0 + v1;
      }
    }
  }

  // ---- Jacobian sparsity (FVMCCSparsity.cxx:47-106): diagonal + one block per interior-face neighbour;
  // columns sorted ascending (PETSc BAIJ keeps rows sorted)
  M.bsrRowPtr.assign(1, 0);
  for (int e = 0; e < nCells; ++e) {
    std::vector<int> cols(1, e);
    for (int k = M.cellFacePtr[e]; k < M.cellFacePtr[e+1]; ++k) {
      const int lf = M.cellFaces[k];
      if (lf < nInner) cols.push_back(M.faceL[lf] == e ? M.faceR[lf] : M.faceL[lf]);
    }
    std::sort(cols.begin(), cols.end());
    M.bsrColInd.insert(M.bsrColInd.end(), cols.begin(), cols.end());
    M.bsrRowPtr.push_back((int)M.bsrColInd.size());
  }
}

// ---- the other ComputeStencil strategies selected by `Setup1.stencil` (a3).  kind: 1 "FaceVertex"
// (ComputeFaceVertexNeighbors.cxx:45-129: every cell sharing a node, no ghosts), 2 "FaceEdge"
// (ComputeFaceEdgeNeighbors.cxx:44-150: cells of HIGHER ID with at least two nodes in common), 3 "Face"
// (ComputeFaceNeighbors.cxx:46-108: interior-face neighbours of higher ID, in cellFaces order).
// CANONICAL ORDER for 1 and 2: the reference walks a std::set<State*> (pointer order of separately allocated
// states, i.e. implementation-defined; allocation order in practice) - here ascending cell ID.
// The one-directional lists of 2 and 3 are what the reference stores: its least-squares loop adds every edge to both
// of its cells (LeastSquareP1PolyRec2D.cxx:100-134), so each edge is listed once.
inline int rebuildStencil(MeshData& M, int kind) {
  if (kind < 1 || kind > 3) return 1;
  const int nCells = M.nCells, nInner = M.nInner;
  std::multimap<int, int> mapNodeState;                       // key = node, value = cell (sorted by key)
  for (int e = 0; e < nCells; ++e)
    for (int k = M.cellNodePtr[e]; k < M.cellNodePtr[e+1]; ++k) mapNodeState.insert(std::make_pair(M.cellNodes[k], e));
  M.stencilPtr.assign(1, 0);
  M.stencil.clear();
  for (int e = 0; e < nCells; ++e) {
    if (kind == 3) {
      for (int k = M.cellFacePtr[e]; k < M.cellFacePtr[e+1]; ++k) {
        const int lf = M.cellFaces[k];
        if (lf >= nInner) continue;                            // "neglect the neighbor state if it is a ghost one"
        const int neighbor = M.faceR[lf];      // state 1 of the face on both branches (:76-88): the cell itself when it is the R cell
        if (neighbor > e) M.stencil.push_back(neighbor);
      }
    } else {
      std::set<int> neighborList;
      const int nbNodesInCell = M.cellNodePtr[e+1] - M.cellNodePtr[e];
      for (int iNode = 0; iNode < nbNodesInCell; ++iNode) {
        const int nodeID = M.cellNodes[M.cellNodePtr[e] + iNode];
        auto range = mapNodeState.equal_range(nodeID);
        for (auto it = range.first; it != range.second; ++it) {
          const int s = it->second;
          if (s == e) continue;
          if (kind == 1) { neighborList.insert(s); continue; }
          if (!(s > e)) continue;
          const int nbNeighborNodes = M.cellNodePtr[s+1] - M.cellNodePtr[s];
          int matchingNodes = 0;
          for (int in = 0; in < nbNodesInCell; ++in)
            for (int ik = 0; ik < nbNeighborNodes; ++ik)
              if (M.cellNodes[M.cellNodePtr[s] + ik] == M.cellNodes[M.cellNodePtr[e] + in]) { ++matchingNodes; break; }
          if (matchingNodes >= 2) neighborList.insert(s);
        }
      }
      M.stencil.insert(M.stencil.end(), neighborList.begin(), neighborList.end());
    }
    M.stencilPtr.push_back((int)M.stencil.size());
  }
  return 0;
}

} // namespace cfo
