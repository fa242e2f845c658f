// COOLFluiD plugin module "FiniteVolumeB200": registers the B200-native ComputeRHS commands with the
// CellCenterFVM space method.  Same registration pattern as plugins/FiniteVolumeCUDA/FiniteVolumeCUDA.hh.
// NOT compiled in this repository's CI (COOLFluiD needs Boost/MPI/PETSc); see INTEGRATION.md.
#ifndef COOLFluiD_Numerics_FiniteVolume_FiniteVolumeB200_hh
#define COOLFluiD_Numerics_FiniteVolume_FiniteVolumeB200_hh

#include "Environment/ModuleRegister.hh"
#include "Common/ExportAPI.hh"

#ifdef FiniteVolumeB200_EXPORTS
#   define FiniteVolumeB200_API CF_EXPORT_API
#else
#   define FiniteVolumeB200_API CF_IMPORT_API
#endif

namespace COOLFluiD { namespace Numerics { namespace FiniteVolume {

class FiniteVolumeB200Module : public Environment::ModuleRegister<FiniteVolumeB200Module> {
public:
  static std::string getModuleName() { return "FiniteVolumeB200"; }
  static std::string getModuleDescription() {
    return "B200-native (sm_100a) FVMCC ComputeRHS / ComputeRhsJacob through libcfb200 (C ABI).";
  }
};

} } }
#endif
