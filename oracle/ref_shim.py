"""Loads the UNMODIFIED reference hot path for differential testing and as the timed CPU baseline.

Two sources, in this order: /root/reference (authoring container: sources read where they lie, patched in
memory) or oracle/_ref/ (staged, already patched, by oracle/make_ref.py; git-ignored, travels to the GPU box,
which has no /root/reference).  tests/golden/ carries the pin where neither exists.

Test infrastructure — never imported by the product.  Three py3/CPU incompatibilities and one print are
patched, and the modules are executed under their own names:
  * models/MA_attention_8_NEW.py:41   hard `.cuda()` on the position embedding -> follow `feats`
  * models/MA_attention_8_NEW.py:125,133  py2 integer `/` in `.view(...)`      -> `//`
  * CaptionGenerator.py:43            per-call `print` of timings               -> dropped
"""
import importlib
import os
import sys
import types

STAGED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
REF_ROOT = os.environ.get("ECHR_REFERENCE_ROOT", "/root/reference")
if not os.path.isfile(os.path.join(REF_ROOT, "CaptionGenerator.py")) and os.path.isfile(os.path.join(STAGED_ROOT, "PATCHED")):
    REF_ROOT = STAGED_ROOT
PREPATCHED = os.path.isfile(os.path.join(REF_ROOT, "PATCHED"))


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "CaptionGenerator.py"))


def source():
    """'checkout' (sources under /root/reference), 'staged' (oracle/_ref) or None."""
    if not available():
        return None
    return "staged" if PREPATCHED else "checkout"


_PATCHES = {
    "models/MA_attention_8_NEW.py": [
        ("requires_grad=False).cuda()", "requires_grad=False).to(feats.device)"),
        ("self.d_q / self.group)", "self.d_q // self.group)"),
        ("self.d_k / self.group).transpose", "self.d_k // self.group).transpose"),
    ],
    "CaptionGenerator.py": [
        ("print(t1 - t0, t2 - t1)", "pass"),
    ],
}

_loaded = None


def _exec_patched(modname, relpath, package=None):
    path = os.path.join(REF_ROOT, relpath)
    src = open(path).read()
    for old, new in ([] if PREPATCHED else _PATCHES.get(relpath, [])):
        if old not in src:
            raise RuntimeError("reference changed: %r not found in %s" % (old, relpath))
        src = src.replace(old, new)
    mod = types.ModuleType(modname)
    mod.__file__ = path
    if package is not None:
        mod.__package__ = package
    sys.modules[modname] = mod
    exec(compile(src, path, "exec"), mod.__dict__)
    return mod


def load():
    """Returns a namespace with the reference classes: CaptionGenerator, MA_Attention8, models."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference not present at %s" % REF_ROOT)
    import warnings
    warnings.filterwarnings("ignore")
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)  # for `misc.utils`
    pkg = types.ModuleType("models")
    pkg.__path__ = [os.path.join(REF_ROOT, "models")]
    pkg.__file__ = os.path.join(REF_ROOT, "models", "__init__.py")
    sys.modules["models"] = pkg
    _exec_patched("models.MA_attention_8_NEW", "models/MA_attention_8_NEW.py", "models")
    importlib.import_module("models.OldModel_NEW")
    importlib.import_module("models.sst_model")
    exec(compile(open(pkg.__file__).read(), pkg.__file__, "exec"), pkg.__dict__)
    cg = _exec_patched("CaptionGenerator", "CaptionGenerator.py")
    _loaded = types.SimpleNamespace(CaptionGenerator=cg.CaptionGenerator, models=pkg,
                                    MA_Attention8=pkg.MA_Attention8)
    return _loaded


def build_reference(opt, state_dict):
    """Reference CaptionGenerator(opt) in eval mode with `state_dict` loaded strictly (eval.py:65-67)."""
    ref = load()
    model = ref.CaptionGenerator(opt.copy() if hasattr(opt, "copy") else opt)
    model.load_state_dict(state_dict, strict=True)
    model.eval()
    return model
