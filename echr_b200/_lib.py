"""ctypes binding of libechr_b200.so (the sm_100a CUDA library behind include/echr_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, an exception
is raised.  torch is only used by callers for device memory and streams; this module passes raw
pointers.
"""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libechr_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "echr_b200.h")


class EchrError(RuntimeError):
    pass


_lib = None

_CTYPES = {
    "int": ctypes.c_int,
    "int64_t": ctypes.c_int64,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
    "const double*": ctypes.c_void_p,
    "const int64_t*": ctypes.c_void_p,
    "void*": ctypes.c_void_p,
    "const void*": ctypes.c_void_p,
    "float*": ctypes.c_void_p,
    "const float*": ctypes.c_void_p,
    "int32_t*": ctypes.c_void_p,
    "int64_t*": ctypes.c_void_p,
    "const int32_t*": ctypes.c_void_p,
    "const char*": ctypes.c_char_p,
    "echr_model_t*": ctypes.c_void_p,
    "echr_model_t**": ctypes.c_void_p,
    "const echr_model_t*": ctypes.c_void_p,
    "echr_batch_t*": ctypes.c_void_p,
    "echr_batch_t**": ctypes.c_void_p,
    "const echr_batch_t*": ctypes.c_void_p,
    "const echr_config_t*": ctypes.c_void_p,
    "const echr_weights_t*": ctypes.c_void_p,
    "const echr_capacity_t*": ctypes.c_void_p,
    "const echr_batch_desc_t*": ctypes.c_void_p,
    "echr_sst_t*": ctypes.c_void_p,
    "echr_sst_t**": ctypes.c_void_p,
    "const echr_sst_weights_t*": ctypes.c_void_p,
}


class Config(ctypes.Structure):
    _fields_ = [("vocab_size", ctypes.c_int32), ("seq_length", ctypes.c_int32), ("a_terms", ctypes.c_int32),
                ("w_terms", ctypes.c_int32), ("clip_bf16", ctypes.c_int32)]


# field order of echr_weights_t (include/echr_b200.h) -> state_dict key
WEIGHT_FIELDS = [
    ("event_emb_w", "fusion_model.event_emb.weight"), ("event_emb_b", "fusion_model.event_emb.bias"),
    ("pos_fc1_w", "fusion_model.enc_attn.pair_pos_fc1.weight"), ("pos_fc1_b", "fusion_model.enc_attn.pair_pos_fc1.bias"),
    ("pos_fc2_w", "fusion_model.enc_attn.pair_pos_fc2.weight"), ("pos_fc2_b", "fusion_model.enc_attn.pair_pos_fc2.bias"),
    ("query_w", "fusion_model.enc_attn.query_1.weight"), ("query_b", "fusion_model.enc_attn.query_1.bias"),
    ("key_w", "fusion_model.enc_attn.key_1.weight"), ("key_b", "fusion_model.enc_attn.key_1.bias"),
    ("out_w", "fusion_model.enc_attn.linear_out_1.weight"), ("out_b", "fusion_model.enc_attn.linear_out_1.bias"),
    ("embed_w", "lm_model.embed.weight"),
    ("logit_w", "lm_model.logit.weight"), ("logit_b", "lm_model.logit.bias"),
    ("lstm_w_ih0", "lm_model.core.layer0.weight_ih"), ("lstm_w_ih1", "lm_model.core.layer1.weight_ih"),
    ("lstm_w_ih2", "lm_model.core.layer2.weight_ih"),
    ("lstm_w_hh0", "lm_model.core.layer0.weight_hh"), ("lstm_w_hh1", "lm_model.core.layer1.weight_hh"),
    ("lstm_w_hh2", "lm_model.core.layer2.weight_hh"),
    ("lstm_b_ih0", "lm_model.core.layer0.bias_ih"), ("lstm_b_ih1", "lm_model.core.layer1.bias_ih"),
    ("lstm_b_ih2", "lm_model.core.layer2.bias_ih"),
    ("lstm_b_hh0", "lm_model.core.layer0.bias_hh"), ("lstm_b_hh1", "lm_model.core.layer1.bias_hh"),
    ("lstm_b_hh2", "lm_model.core.layer2.bias_hh"),
    ("ctx2att_w", "lm_model.core.attention.ctx2att.weight"), ("ctx2att_b", "lm_model.core.attention.ctx2att.bias"),
    ("h2att_w", "lm_model.core.attention.h2att.weight"), ("h2att_b", "lm_model.core.attention.h2att.bias"),
    ("alpha_w", "lm_model.core.attention.alpha_net.weight"),
]


class Weights(ctypes.Structure):
    _fields_ = [(name, ctypes.c_void_p) for name, _ in WEIGHT_FIELDS]


class Capacity(ctypes.Structure):
    _fields_ = [("max_videos", ctypes.c_int64), ("max_clips", ctypes.c_int64), ("max_events", ctypes.c_int64),
                ("max_pairs", ctypes.c_int64), ("max_rows", ctypes.c_int64)]


class BatchDesc(ctypes.Structure):
    _fields_ = [("n_videos", ctypes.c_int32), ("n_events", ctypes.c_int32),
                ("host_clip_offsets", ctypes.c_void_p), ("host_event_video", ctypes.c_void_p),
                ("host_event_start", ctypes.c_void_p), ("host_event_end", ctypes.c_void_p),
                ("host_event_ind", ctypes.c_void_p)]


def declared_symbols(header_path=HEADER_PATH):
    """Parse include/echr_b200.h: {name: (restype, [argtypes])}. The tests use this to check that
    the library exports exactly what the header declares."""
    text = open(header_path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = {}
    for m in re.finditer(r"ECHR_API\s+([\w\s\*]+?)\s*\b(echr_\w+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        ret = re.sub(r"\s+", " ", ret).replace(" *", "*")
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = re.sub(r"\s+", " ", a.strip())
                # drop the parameter name
                mm = re.match(r"(.*?)(\w+)$", a)
                ty = mm.group(1).strip().replace(" *", "*") if mm else a
                argtypes.append(ty)
        out[name] = (ret, argtypes)
    return out


def load():
    """Load the shared library (once). Raises EchrError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EchrError(
            "libechr_b200.so not found at %s — build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `make -C echr_b200/csrc`. There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (ret, args) in declared_symbols().items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = _CTYPES[ret]
        fn.argtypes = [_CTYPES[a] for a in args]
    _lib = lib
    return lib


def last_error():
    return load().echr_last_error().decode("utf-8", "replace")


def check(rc, what=""):
    if rc != 0:
        raise EchrError("%s failed (%d): %s" % (what or "echr call", rc, last_error()))
