// FVMCC_ComputeRHSB200 — drop-in ComputeRHS command: selects the B200 engine from a .CFcase with
//   Simulator.Modules.Libs = ... libFiniteVolume libFiniteVolumeNavierStokes libFiniteVolumeB200
//   Simulator.SubSystem.CellCenterFVM.ComputeRHS = B200Rhs          (explicit)   | B200RhsJacob (implicit, NumJacob) | B200RhsJacobAnalytic (implicit, NumJacobAnalytic)
//   Simulator.SubSystem.CellCenterFVM.B200RhsJacob.Device = 0        (default: rank % deviceCount)
// It derives from FVMCC_ComputeRHS (as plugins/FiniteVolumeCUDA/FVMCC_ComputeRHSCell.hh:31 does) to inherit the
// sockets and the configured strategies, copies socket data to libcfb200 once in setup(), and in execute() sends
// the states, runs the device evaluation and leaves rhs / updateCoeff valid on the host (callers read them right
// away: UpdateSol.cxx:83, FVMCC_PseudoSteadyTimeRhs.cxx:106).  The block Jacobian is inserted in the LSS matrix.
#ifndef COOLFluiD_Numerics_FiniteVolume_FVMCC_ComputeRHSB200_hh
#define COOLFluiD_Numerics_FiniteVolume_FVMCC_ComputeRHSB200_hh

#include <map>
#include "FiniteVolume/FVMCC_ComputeRHS.hh"
#include "Framework/BlockAccumulator.hh"
#include "cfb200.h"

namespace COOLFluiD { namespace Numerics { namespace FiniteVolume {

class FVMCC_ComputeRHSB200 : public FVMCC_ComputeRHS {
public:
  explicit FVMCC_ComputeRHSB200(const std::string& name);
  virtual ~FVMCC_ComputeRHSB200();
  static void defineConfigOptions(Config::OptionList& options);
  virtual void configure(Config::ConfigArgs& args);
  virtual void setup();
  virtual void unsetup();
  virtual void execute();
  /// FVMCC_ComputeRHS's sockets + the reconstruction stencil (socket "stencil", filled by the setup of LeastSquareP1)
  virtual std::vector<Common::SafePtr<Framework::BaseDataSocketSink> > needsSockets();

protected:
  /// 0 for "B200Rhs", 1 for "B200RhsJacob" (= NumJacob), 2 for "B200RhsJacobAnalytic" (= NumJacobAnalytic)
  virtual int withJacobian() const { return 0; }
  void fillConfig(cfb200_config& c);        ///< static part, once in setup()
  void refreshStep(cfb200_config& c);       ///< per-call scalars (resFactor, iteration, residual, CFL, DT, dynamic options)
  void uploadMesh();
  void refreshDirichletValues();            ///< option RefreshInletValues: re-probe the Dirichlet-type commands (Def functions of time / iteration)
  cfb200_bc translateBC(FVMCC_BC* bc, Framework::GeometricEntity* probe, CFuint nEqs);
  void insertJacobian();
  void check(int rc, const char* what);

  Framework::DataSocketSink<std::vector<Framework::State*> > socket_stencil;
  Framework::DataSocketSink<Framework::State*> socket_pastStates;   ///< non-essential (FVMCC_PseudoSteadyTimeRhs.cxx:55)
  Framework::DataSocketSink<CFint> socket_trsID;   ///< non-essential: TRS of each boundary node (NodalStatesExtrapolator.ci:71,485), read for DistanceBasedGMove
  cfb200_handle* m_h;
  CFint m_device;          ///< option Device (-1: rank % device count, as CudaDeviceManager.cu:58)
  bool m_hostJacobian;     ///< option HostJacobian: copy the BSR back and LSSMatrix::addValues (CPU PETSc); else device view
  CFreal m_viscRef, m_viscTRef, m_viscSuthConst, m_fixedVisc;   ///< constants of the viscosity law (verified in setup())
  bool m_fuseTimeRhs, m_timeAnalyticalMatrix, m_timeUseGlobalDT;
  bool m_refreshInletValues;                ///< option RefreshInletValues
  std::vector<std::pair<CFuint, CFint> > m_dirichletTrs;   ///< (TRS index in the TRS list, cfb200_bc record) of the Dirichlet-type commands
  std::vector<bool> m_timeZeroDiag;
  cfb200_config m_cfg;     ///< translated once in setup(), per-call scalars refreshed in execute()
  std::vector<CFreal> m_bsr, m_past;
  std::vector<CFint> m_rowPtr, m_colInd;
  std::map<CFint, Framework::BlockAccumulator*> m_rowAcc;       ///< one 1 x n accumulator per block-row length
};

class FVMCC_ComputeRhsJacobB200 : public FVMCC_ComputeRHSB200 {
public:
  explicit FVMCC_ComputeRhsJacobB200(const std::string& name) : FVMCC_ComputeRHSB200(name) {}
protected:
  virtual int withJacobian() const { return 1; }
};

/// drop-in for FVMCC_ComputeRhsJacobAnalytic ("NumJacobAnalytic", FVMCC_ComputeRhsJacobAnalytic.cxx:32-35)
class FVMCC_ComputeRhsJacobAnalyticB200 : public FVMCC_ComputeRHSB200 {
public:
  explicit FVMCC_ComputeRhsJacobAnalyticB200(const std::string& name) : FVMCC_ComputeRHSB200(name) {}
protected:
  virtual int withJacobian() const { return 2; }
};

} } }
#endif
