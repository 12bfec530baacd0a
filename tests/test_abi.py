"""CPU tests of the drop-in boundary: the shared library loads without a GPU, exports exactly the
symbols include/echr_b200.h declares, and the host-side mirror has the reference's checkpoint layout."""
import ctypes
import subprocess

import pytest
import torch

import helpers
from echr_b200 import _lib
from echr_b200.caption_generator import CaptionGenerator, trim_greedy
from echr_b200.config import echr_opt, validate
from echr_b200.engine import BatchLayout


def test_library_exports_every_declared_symbol():
    decl = _lib.declared_symbols()
    assert len(decl) >= 18
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in decl:
        assert hasattr(lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l and l.split()[-1].startswith("echr_")}
    assert exported == set(decl), exported ^ set(decl)
    assert _lib.load().echr_abi_version() == 5


def test_weights_struct_matches_header():
    text = open(_lib.HEADER_PATH).read()
    body = text[text.index("typedef struct echr_weights {"):text.index("} echr_weights_t;")]
    n_ptr = body.count("const float*")
    n_ptr += 2 * body.count("[3]")  # each [3] array holds three pointers
    assert n_ptr == len(_lib.WEIGHT_FIELDS) == ctypes.sizeof(_lib.Weights) // 8


def test_no_cpu_path():
    m = CaptionGenerator(echr_opt())
    m.load_state_dict(helpers.state_dict("fp32"), strict=True)
    v = helpers.case_video("cfg1")
    with pytest.raises(RuntimeError):
        m(*helpers.video_tuple(v)[:3], [], v.ind_select_list, v.soi_select_list, mode="eval")


def test_unsupported_recipe_is_refused():
    with pytest.raises(ValueError):
        validate(echr_opt(event_context_type="ER1"))
    with pytest.raises(ValueError):
        validate(echr_opt(caption_model="show_attend_tell"))


def test_batch_layout_and_trim():
    lay = BatchLayout([100, 50], [[9, 19], [4]], [[[0, 10], [5, 20]], [[0, 5]]])
    assert lay.n_events == 3 and lay.n_clips == 150 and lay.n_pairs == 5
    assert list(lay.clip_offsets) == [0, 100, 150] and list(lay.event_video) == [0, 0, 1]
    tok = torch.tensor([[5, 0, 0, 0], [7, 8, 0, 0]], dtype=torch.int32)
    lp = torch.zeros(2, 4)
    s, l = trim_greedy(tok, lp)
    assert s.shape == (2, 2) and s.dtype == torch.int64
    assert trim_greedy(torch.zeros(2, 4, dtype=torch.int32), lp) == ([], [])


def test_position_helpers_match_oracle():
    import numpy as np
    from echr_b200.caption_generator import MA_Attention8
    from oracle import echr_oracle as O
    soi = [[0, 10], [5, 6], [20, 60], [3, 99]]
    a = MA_Attention8.extract_position_matrix(np.array(soi), len(soi))
    assert np.array_equal(a, O.position_matrix(soi))
    assert np.array_equal(MA_Attention8.extract_position_embedding(a, 512), O.position_embedding(O.position_matrix(soi)))


def test_submodules_need_their_owner():
    from echr_b200.caption_generator import setup_fusion
    fm = setup_fusion(echr_opt())
    with pytest.raises(RuntimeError):
        fm(torch.zeros(2, 1012), [[0, 1], [1, 2]])


def test_header_is_plain_c99_and_links_from_c(tmp_path):
    """The boundary is a C ABI: the header compiles as strict C99 and a C program links against the library."""
    import os
    import shutil
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    src = tmp_path / "abi.c"
    src.write_text('#include "echr_b200.h"\n'
                   'int main(void) { echr_config_t c; echr_sst_weights_t w; (void)c; (void)w;\n'
                   '  return echr_abi_version() == ECHR_B200_ABI_VERSION ? 0 : 1; }\n')
    inc = os.path.dirname(_lib.HEADER_PATH)
    exe = tmp_path / "abi"
    libdir = os.path.dirname(_lib.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, str(src), "-o", str(exe),
                           "-L", libdir, "-l:libechr_b200.so", "-Wl,-rpath," + libdir])
    assert subprocess.run([str(exe)]).returncode == 0
