"""Import the UNMODIFIED reference from /root/reference.  TEST INFRASTRUCTURE ONLY; build container only.

The GPU box has no /root/reference, so nothing in `-m gpu` tests, smoke() or bench.py goes through here.
Three imports of the reference are unavailable offline and are stubbed:
  kornia      -> oracle.kornia_restated (3 functions; parity unpinned, see that file)
  matplotlib  -> empty module with use()/pyplot
  ipdb        -> empty module with post_mortem()/iex
The reference hard-codes device 'cuda' (optical_flow/raft.py:67,98,116,120; tracker/YAOF...:24,33,95);
`cpu_redirect()` maps those to CPU for the duration of a with-block.
"""
import contextlib
import sys
import types
from pathlib import Path

import torch

REF_ROOT = Path("/root/reference")


def available():
    return (REF_ROOT / "pytracking").is_dir()


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    sys.modules[name] = mod
    return mod


def install_stubs():
    from . import kornia_restated as kr

    if "kornia" not in sys.modules:
        epipolar = _stub("kornia.geometry.epipolar", normalize_points=kr.normalize_points)
        conversions = _stub("kornia.geometry.conversions",
                            convert_points_to_homogeneous=kr.convert_points_to_homogeneous,
                            convert_points_from_homogeneous=kr.convert_points_from_homogeneous)
        homography = _stub("kornia.geometry.homography")
        geometry = _stub("kornia.geometry", epipolar=epipolar, conversions=conversions, homography=homography)
        _stub("kornia", geometry=geometry)
    if "matplotlib" not in sys.modules:
        pyplot = _stub("matplotlib.pyplot")
        cm = _stub("matplotlib.cm")
        _stub("matplotlib", use=lambda *a, **k: None, pyplot=pyplot, cm=cm)
    if "ipdb" not in sys.modules:
        _stub("ipdb", post_mortem=lambda *a, **k: None, iex=lambda f: f, set_trace=lambda *a, **k: None)


def import_reference():
    """Returns a namespace with the reference modules on the hot path."""
    if not available():
        raise RuntimeError("/root/reference is not present (it only exists in the build container)")
    install_stubs()
    for p in (str(REF_ROOT), str(REF_ROOT / "pytracking" / "external" / "RAFT")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from raft_core import corr as ref_corr
        from raft_core import weighted_raft as ref_wraft
        from raft_core.utils import utils as ref_utils
        from pytracking.utils import least_squares_H as ref_lsq
        from pytracking.optical_flow import raft as ref_flow
        from pytracking.utils import misc as ref_misc
        from pytracking.utils import geom_utils as ref_geom
        from pytracking.utils import config as ref_config
    return types.SimpleNamespace(corr=ref_corr, weighted_raft=ref_wraft, utils=ref_utils, lsq=ref_lsq,
                                 flow=ref_flow, misc=ref_misc, geom=ref_geom, config=ref_config)


def irls_without_cuda_check(ref_lsq):
    """`find_homography_IRLSq_QR` refuses CPU tensors (least_squares_H.py:292-293).  Returns a copy of the
    function compiled from the reference's own source with that single check removed, for CPU runs."""
    import inspect
    src = inspect.getsource(ref_lsq.find_homography_IRLSq_QR)
    needle = "    if not points1.is_cuda:\n        raise AssertionError(\"correspondences should be on GPU\")\n"
    assert needle in src, "reference source changed; cannot strip the CUDA check"
    ns = dict(ref_lsq.__dict__)
    exec(compile(src.replace(needle, ""), "<ref IRLSq_QR without is_cuda check>", "exec"), ns)
    return ns["find_homography_IRLSq_QR"]


@contextlib.contextmanager
def cpu_redirect():
    """Within the block, `.to('cuda')` on tensors and modules is a no-op move to CPU."""
    t_to, m_to = torch.Tensor.to, torch.nn.Module.to

    def _fix(args, kwargs):
        args = tuple("cpu" if (isinstance(a, str) and a.startswith("cuda")) else a for a in args)
        kwargs = {k: ("cpu" if (isinstance(v, str) and v.startswith("cuda")) else v) for k, v in kwargs.items()}
        return args, kwargs

    def tensor_to(self, *args, **kwargs):
        args, kwargs = _fix(args, kwargs)
        return t_to(self, *args, **kwargs)

    def module_to(self, *args, **kwargs):
        args, kwargs = _fix(args, kwargs)
        return m_to(self, *args, **kwargs)

    torch.Tensor.to, torch.nn.Module.to = tensor_to, module_to
    try:
        yield
    finally:
        torch.Tensor.to, torch.nn.Module.to = t_to, m_to
