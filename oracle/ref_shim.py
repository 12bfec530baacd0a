"""Loads the UNMODIFIED reference hot path from /root/reference for differential testing (authoring
container only; the GPU box has no /root/reference and uses tests/golden/ instead).

Test infrastructure — never imported by the product.  Nothing is copied: the reference sources are
read where they lie, three py3/CPU incompatibilities are patched in memory, and the modules are
executed under their own names:
  * models/MA_attention_8_NEW.py:41   hard `.cuda()` on the position embedding -> follow `feats`
  * models/MA_attention_8_NEW.py:125,133  py2 integer `/` in `.view(...)`      -> `//`
  * CaptionGenerator.py:43            per-call `print` of timings               -> dropped
"""
import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("ECHR_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "CaptionGenerator.py"))


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
    for old, new in _PATCHES.get(relpath, []):
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
