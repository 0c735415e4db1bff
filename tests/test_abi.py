"""The C-ABI library loads and exports every symbol include/pcb.h declares (no compute calls: CPU only)."""
import ctypes as C
import re

from pcutil import ROOT, load_pcb


def test_every_declared_symbol_is_exported():
    header = (ROOT / "include" / "pcb.h").read_text()
    declared = sorted(set(re.findall(r"\b(pcb_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 20
    pcb = load_pcb()
    lib = pcb.load_library()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in pcb.h but not exported by libpcb.so"
    assert sorted(pcb.EXPORTS) == declared, "pcb.py EXPORTS out of sync with pcb.h"


def test_struct_sizes_match_header():
    pcb = load_pcb()
    assert C.sizeof(pcb.PcbOptions) == 32
    assert C.sizeof(pcb.PcbNodeInfo) == 32
    assert C.sizeof(pcb.PcbKernelStat) == 64
    assert pcb.POINT_DTYPE.itemsize == 16


def test_no_device_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        return
    pcb = load_pcb()
    try:
        pcb.Converter(None)
    except pcb.PcbError as e:
        assert e.status == -4   # PCB_ERR_CUDA: there is no CPU path
    else:
        raise AssertionError("pcb_open succeeded without a CUDA device")
