"""Model flags and the SoA parameter-block layout (host mirror of include/kima_cuda.h).

The nine flags are the compile-time constants of a ``kima_setup.cpp``
(/root/reference/src/RVmodel.h:16-34, src/RVConditionalPrior.h:11) plus ``npmax``
from the ``RVmodel`` constructor (e.g. examples/51Peg/kima_setup.cpp:13).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, fields


class CModelFlags(ctypes.Structure):
    """``kc_model_flags`` of include/kima_cuda.h."""

    _fields_ = [(n, ctypes.c_int32) for n in (
        "GP", "MA", "hyperpriors", "trend", "degree", "multi_instrument",
        "known_object", "n_known_object", "studentt", "npmax", "kernel")]


class CLayout(ctypes.Structure):
    """``kc_layout`` of include/kima_cuda.h."""

    _fields_ = [(n, ctypes.c_int32) for n in (
        "n_rows", "npmax", "row_P", "row_K", "row_phi", "row_ecc", "row_w",
        "row_vsys", "n_jit", "row_jit", "n_off", "row_off", "row_trend",
        "n_beta", "row_beta", "row_eta", "row_nu", "n_ko", "row_KO_P",
        "row_KO_K", "row_KO_e", "row_KO_phi", "row_KO_w", "row_MA")]

    def as_dict(self) -> dict:
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


@dataclass(frozen=True)
class ModelFlags:
    """Same names and meaning as the reference's ``extern const`` flags."""

    GP: bool = False
    MA: bool = False
    hyperpriors: bool = False
    trend: bool = False
    degree: int = 0
    multi_instrument: bool = False
    known_object: bool = False
    n_known_object: int = 0
    studentt: bool = False
    npmax: int = 1
    kernel: int = 0  # 0 = standard quasi-periodic (src/RVmodel.h:63-64)

    def __post_init__(self):
        # Same setup errors as RVmodel::setPriors (src/RVmodel.cpp:43-47).
        if self.trend and self.degree == 0:
            raise ValueError("trend=true but degree=0, what gives?")
        if self.trend and self.degree > 3:
            raise ValueError("can't go higher than 3rd degree trends")
        if self.npmax < 0:
            raise ValueError("npmax must be >= 0")

    def to_c(self) -> CModelFlags:
        c = CModelFlags()
        for f in fields(self):
            setattr(c, f.name, int(getattr(self, f.name)))
        return c


def layout_from_flags(flags: ModelFlags, n_inst: int, n_ind: int) -> dict:
    """Python restatement of ``kc_layout_from_flags`` (include/kima_cuda.h).

    Kept only so host code can reason about the block without a library
    handle; tests pin it against the C function.
    """
    r = 0
    L = {"npmax": flags.npmax}
    for name in ("row_P", "row_K", "row_phi", "row_ecc", "row_w"):
        L[name] = r
        r += flags.npmax
    L["row_vsys"] = r
    r += 1
    L["n_jit"] = n_inst if flags.multi_instrument else 1
    L["row_jit"] = r
    r += L["n_jit"]
    L["n_off"] = (n_inst - 1) if flags.multi_instrument else 0
    L["row_off"] = r if L["n_off"] else -1
    r += L["n_off"]
    L["row_trend"] = r if flags.trend else -1
    r += 3 if flags.trend else 0
    L["n_beta"] = n_ind
    L["row_beta"] = r if n_ind else -1
    r += n_ind
    L["row_eta"] = r if flags.GP else -1
    r += 4 if flags.GP else 0
    L["row_nu"] = r if flags.studentt else -1
    r += 1 if flags.studentt else 0
    L["n_ko"] = flags.n_known_object if flags.known_object else 0
    for name in ("row_KO_P", "row_KO_K", "row_KO_e", "row_KO_phi", "row_KO_w"):
        L[name] = r if L["n_ko"] else -1
        r += L["n_ko"]
    L["row_MA"] = r if flags.MA else -1
    r += 2 if flags.MA else 0
    L["n_rows"] = r
    return L
