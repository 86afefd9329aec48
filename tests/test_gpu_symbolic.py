"""GPU test of the device-side symbolic helpers: the update-pair list and the relative maps
built on the GPU from L->super / L->pi / L->s must be bit-identical to the host planner's tables, and a
factorization that uses them (the default for a one-GPU context) must match the reference."""
import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import b200
from cases import get_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,size,ordering", [("lap2d5", 100, refd.CHOLMOD_AMD), ("lap3d7", 30, refd.CHOLMOD_METIS),
                                                ("lap2d9", 150, refd.CHOLMOD_METIS), ("el27", 10, refd.CHOLMOD_METIS)])
def test_device_pairs_and_relative_maps_equal_the_host_planners(kind, size, ordering):
    c = get_case(kind, size, ordering)
    H = b200.Factor(c.sym, device=-1)                   # host planner: builds the relative maps itself
    F = b200.Factor(c.sym, device=0)                    # one-GPU context: builds them on the device
    assert F.geti("device_symbolic") == 1 and F.getd("ms_symbolic") > 0
    assert F.geti("npairs") == H.geti("npairs") and F.geti("relmap_len") == H.geti("relmap_len") > 0
    assert np.array_equal(F.plan_table("dev_pairs", b200.PAIR_DT), H.plan_table("pairs", b200.PAIR_DT))
    assert np.array_equal(F.plan_table("dev_pair_map", np.int64), H.plan_table("pair_map", np.int64))
    assert np.array_equal(F.plan_table("dev_relmap", np.int32), H.plan_table("relmap", np.int32))
    # ... and the host side of the device context never built them
    assert F.plan_table("relmap", np.int32).size == 0
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == b200.OK and c.rel_err(Lx) <= 1e-10
    F.close()
    H.close()


def test_host_relative_maps_on_request(monkeypatch):
    c = get_case("lap3d7", 16, refd.CHOLMOD_METIS)
    monkeypatch.setenv("CHOLMOD_B200_HOST_RELMAP", "1")
    F = b200.Factor(c.sym, device=0)
    assert F.geti("device_symbolic") == 0 and F.plan_table("relmap", np.int32).size == F.geti("relmap_len")
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == b200.OK and c.rel_err(Lx) <= 1e-10
    F.close()
