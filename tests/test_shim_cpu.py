"""The drop-in boundary as a compiled artefact: libyolo_core_b200.so exports every entry point a yolo-core build binds on
this path, with the reference's names (NmsSort: the reference's C++-mangled name), and was compiled against struct
layouts identical to the ones the reference itself was compiled with (oracle/_ref)."""
import ctypes as C

import pytest

from darknet_b200 import shim
from oracle import pyoracle

pytestmark = pytest.mark.skipif(not shim.have(), reason="libyolo_core_b200.so not built (needs the reference header tree)")


def test_shim_exports_reference_entry_points(built):
    L = shim.lib()
    missing = [s for s in shim.REFERENCE_SYMBOLS if not hasattr(L, s)]
    assert not missing, missing


@pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built")
def test_shim_struct_layouts_match_the_compiled_reference(built):
    L, R = shim.lib(), pyoracle.ref_lib()
    for what, name in enumerate(["layer", "Network", "NetworkState", "Detection"]):
        assert L.yc_shim_sizeof(what) == R.ref_sizeof(what), name
    assert C.sizeof(shim.RefDetection) == L.yc_shim_sizeof(3)
