"""The C-ABI library loads on a CPU-only box and exports every symbol include/cor_b200.h declares."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "cor_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cor_[a-z_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    from cor_asv_ann_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), n
        assert n in _lib.PROTOTYPES, "no ctypes prototype for " + n
    assert lib.cor_abi_version() == 2


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    from cor_asv_ann_b200.engine import Engine
    with pytest.raises(RuntimeError):
        Engine(dict(width=32, depth=1, voc_size=8), {})


def test_library_holds_blackwell_instructions():
    """The built library is sm_100a code that uses the 5th-generation tensor cores and TMA: `cuobjdump -sass` shows
    tcgen05.mma (UTCHMMA, also as CTA pairs), TMA tensor loads and stores (UTMALDG / UTMASTG), TMEM loads (LDTM), bulk
    copies (UBLKCP) and the tcgen05 commit (UTCBAR) -- the mnemonics /opt/skills/guides/B200_PROFILING.md names."""
    import shutil
    import subprocess
    from cor_asv_ann_b200 import _lib
    if shutil.which("cuobjdump") is None:
        pytest.skip("no cuobjdump on this machine")
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    out = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True, timeout=300).stdout
    assert "sm_100a" in out
    for mnemonic in ("UTCHMMA", "UTCHMMA.2CTA", "UTMALDG.2D", "UTMASTG.2D", "LDTM", "UBLKCP", "UTCBAR"):
        assert mnemonic in out, mnemonic
    for kernel in ("gemm_tc_kernel", "attention_tma_kernel", "beam_push_kernel", "page_decoder_kernel", "tr_cell_fwd"):
        assert kernel in out, kernel
