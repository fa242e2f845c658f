"""GPU tier, N > 1 (needs >= 2 B200s: `gpurun --gpus 2`; skipped on a single-GPU box): one process per GPU, each with
its partition on cuda:<rank>; states travel through cfb200_halo_begin / cfb200_halo_exchange (NCCL over NVLink, every
other exchange overlapped with the next ComputeRHS), the residual norm through cfb200_allreduce_sum.  Same bar as the
gloo test of the host logic: updatable + overlap states equal the single-GPU run, residual history to 1e-10."""
import pytest

from test_halo_gloo import launch, check_against_single_domain

pytestmark = pytest.mark.gpu


def _ngpu(product_lib):
    return product_lib.device_count()


@pytest.mark.parametrize("name,world,renumber", [("ns3d_hex", 2, 0), ("euler2d_2nd_tri_fullstencil", 2, 0), ("ns3d_tet", 4, 0),
                                                 ("ns3d_hex", 2, 1)])
def test_forward_euler_with_halo_nccl(name, world, renumber, tmp_path, product_lib, oracle):
    if _ngpu(product_lib) < world:
        pytest.skip("needs %d GPUs" % world)
    steps = 6
    launch(world, "cuda", name, steps, tmp_path, renumber=renumber)
    check_against_single_domain(world, name, steps, tmp_path, product_lib)
