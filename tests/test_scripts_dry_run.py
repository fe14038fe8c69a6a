"""Dry runs of the GPU check scripts (scripts/limiter_gpu_check.py, scripts/eos_tables_gpu_check.py) on the CPU
with a MOCK of HydroBackend that answers from the checkers (the C oracle; the compiled reference where it is built):
they exercise the scripts' own logic -- fixture keys, indexing, the per-step comparisons, the census of the log --
so that a Python-level mistake there does not wait for a GPU to show.  They say NOTHING about the device code: the
mock is test infrastructure and exists in this file only."""
import os
import sys

import numpy as np
import pytest

import integ_lib as il
import oracle_lib as ol
import parity

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


class LimiterMock:
    def __init__(self, prm, device=0):
        self.prm = prm; self.sx = None; self.n = 0
    def sphenix_set_props(self, **kw): self.sx = dict(kw)
    def build_from_cells(self, ic, cells, tasks, nt):
        self.ic = {k: (np.array(v) if v is not None else None) for k, v in ic.items()}; self.tree = (cells, tasks, nt); self.n = len(ic["h"])
    def integrator_init(self, **ip):
        self.ip = ip
        self.kw = dict(periodic=bool(self.prm.periodic), scheme=int(self.prm.scheme), alpha=float(self.prm.viscosity_alpha))
        self.s = il.new_state(self.ic); n = self.n
        self.s["f"] = np.zeros(n, np.float32); self.s["balsara"] = np.zeros(n, np.float32)
        self.variant = {0: "minimal", 1: "planetary", 2: "sphenix"}[int(self.prm.scheme)]
        if self.sx is not None:
            self.sxp = dict(ol.SPHENIX_PROPS, time_base=ip["time_base"])
            self.s["alpha_visc"] = np.full(n, self.kw["alpha"], np.float32)
            self.s["alpha_diff"] = np.full(n, self.sxp.get("diffusion_alpha", 0.0), np.float32)
            for k in ("div_v_prev", "div_v_dt", "laplace_u"): self.s[k] = np.zeros(n, np.float32)
        else: self.sxp = None
        self.chk = il.Checker("oracle", ip, variant=self.variant)
        self.mass = np.array(self.ic["mass"], np.float32); self.ids = np.arange(n, dtype=np.int64); self.h_build = self.s["h"].copy()
    def set_time(self, ti, mab): self.ti, self.mab = ti, mab
    def drift(self, ti_old): return self.chk.drift(self.s, ti_old, self.ti)
    def step(self, do_sort=True):
        il.hydro_step("oracle", self.s, self.kw, self.tree, self.mab, variant=self.variant, mass=self.mass, ids=self.ids, h_build=self.h_build, sphenix=self.sxp)
    def kick(self, w): self.chk.kick(self.s, w, self.ti, self.mab)
    def timestep(self): return self.chk.timestep(self.s, self.ti, self.mab)
    def state(self): return {k: self.s[k].copy() for k in il.STATE_KEYS}
    def sphenix_results(self):
        return {"alpha_visc": self.s["alpha_visc"], "alpha_diff": self.s["alpha_diff"], "div_v_previous_step": self.s["div_v_prev"], "div_v_dt": self.s["div_v_dt"], "laplace_u": self.s["laplace_u"]}
    def runner_do_limiter_loop(self, time_bin=None, want_wakeup=False):
        self.wake = il.limiter_loop("oracle", self.s, self.kw, self.tree, self.mab, variant=self.variant, mass=self.mass, ids=self.ids, h_build=self.h_build, sphenix=self.sxp)
        return self.wake.copy() if want_wakeup else None
    def runner_do_limiter(self): return self.chk.limiter(self.s, self.wake, self.ti, self.mab)
    def limiter_timings(self): return {"limiter_loop_ms": 1.0, "limiter_ms": 1.0}
    def close(self): pass



def test_limiter_check_script_flow_with_a_mock_backend(monkeypatch, capsys):
    import limiter_gpu_check as L
    monkeypatch.setattr(L, "HydroBackend", LimiterMock)
    monkeypatch.setattr(L, "FAILED", [])
    L.run_checks()
    assert L.FAILED == [], capsys.readouterr().out


class EosMock:
    def __init__(self, prm, device=0):
        self.prm = prm; self.lib = __import__("swift_subtask_dev_b200").load_library()
    def set_eos_table(self, t): assert t.kind in (1, 2) and t.P
    def build_from_cells(self, ic, cells, tasks, nt): self.ic, self.tree, self.n = ic, (cells, tasks, nt), len(ic["h"])
    def record_pairs(self, cap): pass
    def _kw(self):
        p = self.prm
        return dict(periodic=bool(p.periodic), dim=tuple(p.dim), scheme=1, alpha=float(p.viscosity_alpha), h_max=float(p.h_max))
    def step(self, do_sort=True):
        if getattr(self, "integ", False):
            il.hydro_step("ref", self.s, self._kw(), self.tree, self.mab, variant="planetary", mass=self.mass, ids=self.ids, h_build=self.h_build)
            return
        cells, tasks, nt = self.tree
        prm = ol.make_params(record=True, **self._kw())
        self.r = ol.run_step("ref", prm, self.ic, cells, tasks, nt, record=True, variant="planetary")
    def results(self, first_pass=False):
        out = {kg: self.r[ko] for kg, ko in parity.FIELD_MAP}
        return out
    def counters(self): return {"n_density": self.r["counts"][0], "n_density_redo": self.r["counts"][1], "n_force": self.r["counts"][2]}
    def fetch_pairs(self, k, cap): return self.r["pairs"][k]
    def integrator_init(self, **ip):
        self.integ = True; self.ip = ip; self.s = il.new_state(self.ic); n = self.n
        self.s["f"] = np.zeros(n, np.float32); self.s["balsara"] = np.zeros(n, np.float32)
        self.chk = il.Checker("ref", ip, variant="planetary")
        self.mass = np.array(self.ic["mass"], np.float32); self.ids = np.array(self.ic["id"], np.int64); self.h_build = self.s["h"].copy()
    def set_time(self, ti, mab): self.ti, self.mab = ti, mab
    def drift(self, ti_old): return self.chk.drift(self.s, ti_old, self.ti)
    def kick(self, w): self.chk.kick(self.s, w, self.ti, self.mab)
    def timestep(self): return self.chk.timestep(self.s, self.ti, self.mab)
    def state(self): return {k: self.s[k].copy() for k in il.STATE_KEYS}
    def close(self): pass


@pytest.mark.skipif(not ol.ref_available("planetary"), reason="oracle/_ref not built (needs the reference tree)")
def test_eos_tables_check_script_flow_with_a_mock_backend(monkeypatch, capsys):
    import eos_tables_gpu_check as E
    import make_golden_eos_step as gen
    gen.load_reference_tables()
    monkeypatch.setattr(E, "HydroBackend", EosMock)
    monkeypatch.setattr(E, "FAILED", [])
    E.step_checks()
    E.run_checks()
    assert E.FAILED == [], capsys.readouterr().out


def test_edge_cases_check_script_flow_with_a_mock_backend(monkeypatch, capsys):
    """scripts/edge_cases_gpu_check.py with the GPU run replaced by the oracle on the tree the library's HOST builder
    makes (swift_cuda_host_plan, no GPU): the script's flow, and that the oracle itself takes one particle, two,
    nobody active, one in a hundred active without aborting."""
    import edge_cases_gpu_check as E
    from swift_subtask_dev_b200 import backend as B

    def mock_run_gpu(ic, kw, record=False, pair_capacity=0, tree=None, eos=None, fields_mode=0):
        t = B.host_plan(parity.gpu_params(kw), ic)
        o = parity.run_oracle_on_tree(ic, t, kw, record=True)
        out = {kg: o[ko] for kg, ko in parity.FIELD_MAP}
        out["counters"] = {"n_density": o["counts"][0], "n_density_redo": o["counts"][1], "n_force": o["counts"][2]}
        out["tree"], out["pairs"] = t, o["pairs"]
        return out

    class EmptyMock:
        def __init__(self, prm):
            pass

        def build(self, ic):
            assert len(ic["h"]) == 0

        def step(self):
            pass

        def counters(self):
            return {"n_density": 0, "n_force": 0}

        def close(self):
            pass

    monkeypatch.setattr(parity, "run_gpu", mock_run_gpu)
    monkeypatch.setattr(E, "HydroBackend", EmptyMock)
    monkeypatch.setattr(E, "FAILED", [])
    assert E.main() == 0, capsys.readouterr().out
