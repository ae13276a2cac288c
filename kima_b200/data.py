"""Host mirror of the reference's ``Data`` singleton (src/Data.h, src/Data.cpp).

Same loaders, argument meaning and derived quantities; the arrays it holds are
exactly what ``kc_upload_data`` sends to the GPU once per data set, replacing the
by-value copy of the singleton that every reference hot-path call makes
(src/RVmodel.cpp:173,288,829).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np


def _read_records(filename: str) -> List[List[float]]:
    """One record per line: the leading run of tokens that parse as doubles.

    Follows ``operator>>(istream&, record_t&)`` / ``operator>>(istream&,
    data_t&)`` (src/Data.cpp:25-69): a header line yields an empty record that
    still counts towards ``skip``.
    """
    try:
        fh = open(filename, "r")
    except OSError:
        # src/Data.cpp:118-122 prints and exit(1)s; raise instead of killing Python
        raise FileNotFoundError("Could not read data file (%s)!" % filename)
    records = []
    with fh:
        for line in fh.read().split("\n"):
            rec = []
            for tok in line.split():
                try:
                    rec.append(float(tok))
                except ValueError:
                    break
            records.append(rec)
    # a trailing newline leaves no extra record (getline fails at EOF)
    if records and records[-1] == []:
        records.pop()
    return records


@dataclass
class RVData:
    """RV time series: t, y, sig (m/s), 1-based instrument ids, indicators."""

    t: np.ndarray = field(default_factory=lambda: np.zeros(0))
    y: np.ndarray = field(default_factory=lambda: np.zeros(0))
    sig: np.ndarray = field(default_factory=lambda: np.zeros(0))
    obsi: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    actind: np.ndarray = field(default_factory=lambda: np.zeros((0, 0)))
    number_instruments: int = 1
    indicator_names: List[str] = field(default_factory=list)
    datamulti: bool = False
    datafile: str = ""
    datafiles: List[str] = field(default_factory=list)
    dataunits: str = "ms"
    dataskip: int = 0

    # ---- constructors mirroring Data::load / load_multi ---------------------
    @classmethod
    def from_arrays(cls, t, y, sig, obsi=None, actind=None, n_inst: Optional[int] = None):
        t = np.ascontiguousarray(t, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        sig = np.ascontiguousarray(sig, dtype=np.float64)
        if not (t.shape == y.shape == sig.shape) or t.ndim != 1:
            raise ValueError("t, y, sig must be 1-D arrays of equal length")
        if obsi is None:
            obsi_a = np.ones(t.size, np.int32)
            ni = 1 if n_inst is None else n_inst
        else:
            obsi_a = np.ascontiguousarray(obsi, dtype=np.int32)
            ni = int(len(set(obsi_a.tolist()))) if n_inst is None else n_inst
            if obsi_a.size and (obsi_a.min() < 1 or obsi_a.max() > ni):
                raise ValueError("obsi must be 1-based instrument ids <= n_inst")
        if actind is None:
            act = np.zeros((0, t.size))
        else:
            act = np.ascontiguousarray(actind, dtype=np.float64).reshape(-1, t.size)
        return cls(t=t, y=y, sig=sig, obsi=obsi_a, actind=act,
                   number_instruments=ni, datamulti=ni > 1)

    @classmethod
    def load(cls, filename: str, units: str = "ms", skip: int = 2,
             indicators: Sequence[str] = ()):
        """``Data::load`` (src/Data.cpp:86-195): one file, one instrument."""
        recs = _read_records(filename)[skip:]
        factor = 1e3 if units == "kms" else 1.0
        names = [n for n in indicators if n != ""]
        cols = [3 + i for i, n in enumerate(indicators) if n != ""]
        t = np.array([r[0] for r in recs], dtype=np.float64)
        y = np.array([r[1] * factor for r in recs], dtype=np.float64)
        sig = np.array([r[2] * factor for r in recs], dtype=np.float64)
        act = np.array([[r[c] * factor for r in recs] for c in cols],
                       dtype=np.float64).reshape(len(cols), t.size)
        # subtract means from activity indicators (src/Data.cpp:166-173)
        for j in range(act.shape[0]):
            mean = math.fsum([]) if t.size == 0 else _accumulate(act[j]) / y.size
            act[j] = act[j] - mean
        d = cls(t=t, y=y, sig=sig, obsi=np.ones(t.size, np.int32), actind=act,
                number_instruments=1, indicator_names=names, datamulti=False,
                datafile=filename, dataunits=units, dataskip=skip)
        return d

    @classmethod
    def load_multi(cls, filename_or_files, units: str = "ms", skip: int = 2,
                   indicators: Sequence[str] = ()):
        """``Data::load_multi``: one joined file with an id column
        (src/Data.cpp:210-284) or several files (src/Data.cpp:301-463)."""
        factor = 1e3 if units == "kms" else 1.0
        if isinstance(filename_or_files, (str, bytes)):
            recs = _read_records(filename_or_files)[skip:]
            # consecutive distinct ids -> 1..n_inst (src/Data.cpp:247-269)
            obsi, idx, last = [], 0, None
            for r in recs:
                if last is not None and r[3] != last:
                    idx += 1
                last = r[3]
                obsi.append(idx + 1)
            t = np.array([r[0] for r in recs], dtype=np.float64)
            y = np.array([r[1] * factor for r in recs], dtype=np.float64)
            sig = np.array([r[2] * factor for r in recs], dtype=np.float64)
            obsi = np.array(obsi, dtype=np.int32)
            return cls(t=t, y=y, sig=sig, obsi=obsi, actind=np.zeros((0, t.size)),
                       number_instruments=len(set(obsi.tolist())), datamulti=True,
                       datafile=filename_or_files, dataunits=units, dataskip=skip)

        files = list(filename_or_files)
        names = [n for n in indicators if n != ""]
        cols = [3 + i for i, n in enumerate(indicators) if n != ""]
        recs, obsi = [], []
        for k, fn in enumerate(files):
            r = _read_records(fn)[skip:]
            recs += r
            obsi += [k + 1] * len(r)
        t = np.array([r[0] for r in recs], dtype=np.float64)
        y = np.array([r[1] * factor for r in recs], dtype=np.float64)
        sig = np.array([r[2] * factor for r in recs], dtype=np.float64)
        obsi = np.array(obsi, dtype=np.int32)
        act = np.array([[r[c] * factor for r in recs] for c in cols],
                       dtype=np.float64).reshape(len(cols), t.size)
        n_inst = len(set(obsi.tolist()))
        if n_inst > 1:
            # sort merged instruments by time (src/Data.cpp:423-458)
            order = np.argsort(t, kind="stable")
            t, y, sig, obsi = t[order], y[order], sig[order], obsi[order]
            act = act[:, order]
        return cls(t=t, y=y, sig=sig, obsi=obsi, actind=np.ascontiguousarray(act),
                   number_instruments=n_inst, indicator_names=names, datamulti=True,
                   datafiles=files, dataunits=units, dataskip=skip)

    # ---- the getters of src/Data.h:52-108 -----------------------------------
    def N(self) -> int:
        return int(self.t.size)

    @property
    def number_indicators(self) -> int:
        return int(self.actind.shape[0])

    @property
    def indicator_correlations(self) -> bool:
        return self.number_indicators > 0

    @property
    def M0_epoch(self) -> float:
        """Epoch of the mean anomaly = first time stamp (src/Data.cpp:162,280,461)."""
        return float(self.t[0])

    def get_t(self): return self.t
    def get_y(self): return self.y
    def get_sig(self): return self.sig
    def get_obsi(self): return self.obsi
    def get_actind(self): return self.actind
    def get_t_min(self): return float(self.t.min())
    def get_t_max(self): return float(self.t.max())
    def get_timespan(self): return self.get_t_max() - self.get_t_min()
    def get_t_middle(self): return self.get_t_min() + 0.5 * self.get_timespan()
    def get_RV_min(self): return float(self.y.min())
    def get_RV_max(self): return float(self.y.max())
    def get_RV_span(self): return self.get_RV_max() - self.get_RV_min()

    def get_max_RV_span(self) -> float:
        """src/Data.cpp:527-556."""
        if not self.datamulti:
            return self.get_RV_span()
        span = 0.0
        for j in range(self.number_instruments):
            yy = self.y[self.obsi == j + 1]
            span = max(span, float(yy.max() - yy.min()))
        return span

    def get_RV_var(self) -> float:
        """src/Data.cpp:466-477 (sample variance, N-1)."""
        mean = _accumulate(self.y) / self.y.size
        return float(_accumulate((self.y - mean) ** 2) / (self.y.size - 1))

    def get_RV_std(self) -> float:
        return math.sqrt(self.get_RV_var())

    def topslope(self) -> float:
        """src/Data.cpp:488-517."""
        if not self.datamulti:
            return self.get_RV_span() / self.get_timespan()
        slope = 0.0
        for j in range(self.number_instruments):
            m = self.obsi == j + 1
            yy, tt = self.y[m], self.t[m]
            slope = max(slope, float((yy.max() - yy.min()) / (tt.max() - tt.min())))
        return slope

    def get_trend_magnitude(self, degree: int) -> int:
        """src/Data.cpp:596-600."""
        return int(round(math.log10(self.get_RV_span() / self.get_timespan() ** degree)))

    def Ninstruments(self) -> int:
        return len(set(self.obsi.tolist()))


def _accumulate(a) -> float:
    """Left-to-right double sum, as ``std::accumulate(..., 0.0)``."""
    s = 0.0
    for v in np.asarray(a, dtype=np.float64).tolist():
        s += v
    return s
