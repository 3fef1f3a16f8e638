// Grid / XC / Fock kernel instantiations of the MGGA functional family (sadatom_grid_mma.cuh).
#include "../../include/helfem_b200.h"
#include "sadatom_grid_launch.cuh"

namespace hfb {
int launch_grid_mgga(int nspin, const SadGridParams& prm, cudaStream_t st) { return launch_grid_family<xc::FAM_MGGA>(nspin, prm, st); }
}  // namespace hfb
