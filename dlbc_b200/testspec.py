"""The reference's JSON test specifications, read and executed against the GPU path -- SURVEY.md 8(f) row 1.

A DLBC regression test is a ``<name>.json`` next to its ``<name>.in`` (tests/runnable/dlbct/test.py:16-124):
input file, an optional parameter matrix (``parameters``: every combination of the listed values is a subtest),
overrides for the quick (``fast``) and the coverage run, and ``compare`` rules -- for each ``data`` stem and each
``comparison`` of type ``h5diff`` a file pattern ``%data%-<sim>-*t<time>.h5`` that must match the file of the same
name under ``reference-data`` (dlbct/compare.py:59-88), exactly or, with ``accuracy``, to that ABSOLUTE difference
(``h5diff -d``; the reference applies it to every compiler but DMD, whose output the goldens are).

`TestSpec` parses such a file; `run` replays one subtest through `dlbc_b200.replay` and applies the compare rules to
the arrays the replay collected, against golden arrays handed in by the caller (the tests read them from
tests/golden/*.npz, converted from the reference's reference-data by tests/golden/make_golden.py).
"""
from __future__ import annotations

import itertools
import json
import os
import re

import numpy as np

from . import params as P
from . import replay as R


class SpecError(ValueError):
    pass


class TestSpec:
    __test__ = False  # (not a pytest class)

    def __init__(self, path):
        self.path = path
        self.root = os.path.dirname(os.path.abspath(path))
        try:
            with open(path) as f:
                self.data = json.load(f)
        except ValueError as e:
            raise SpecError(f"JSON file '{path}' seems to be broken: {e}") from None
        for key in ("name", "description", "configuration", "input-file", "clean", "compare"):  # test.py:33-66
            if key not in self.data:
                raise SpecError(f"JSON file '{path}' lacks a '{key}' parameter.")
        d = self.data
        self.name, self.configuration, self.input_file = d["name"], d["configuration"], d["input-file"]
        self.tags = d.get("tags", [])
        self.np = d.get("np", 1)
        self.checkpoint = d.get("checkpoint")
        self.parameters = d.get("parameters") or []
        self.fast = d.get("fast")
        self.coverage = d.get("coverage")
        self.compare = d["compare"]

    # ---- what to run ------------------------------------------------------------------------------------
    def subtests(self):
        """Every combination of the parameter matrix (test.py:117-118, run.py): list of ["name=value", ...]."""
        if not self.parameters:
            return [[]]
        axes = [[f'{p["parameter"]}={v}' for v in p["values"]] for p in self.parameters]
        return [list(c) for c in itertools.product(*axes)]

    def overrides(self, mode="fast"):
        """--parameter overrides of the quick / coverage run (run.py: fast and coverage parameter sets)."""
        block = {"fast": self.fast, "coverage": self.coverage, "full": None}[mode]
        if not block:
            return []
        return [f'{p["parameter"]}={p["value"]}' for p in block.get("parameters", [])]

    def compare_rules(self, mode="fast"):
        """[(data stem, timestep, absolute accuracy or None)] of the h5diff comparisons that apply to `mode`."""
        block = {"fast": self.fast, "coverage": self.coverage, "full": None}[mode]
        cmp_ = (block or {}).get("compare") or self.compare
        rules = []
        for c in cmp_.get("comparison", []):
            if c.get("type") != "h5diff":
                raise SpecError(f"Unknown comparison type '{c.get('type')}'.")  # compare.py:86
            m = re.search(r"t(\d{8})\.h5$", c["files"])  # ("%data%-<sim>*-t00000005.h5")
            if not m:
                raise SpecError(f"cannot find the timestep in the file pattern '{c['files']}'")
            acc = float(c["accuracy"]) if "accuracy" in c else None
            for d in cmp_.get("data", []):
                rules.append((d, int(m.group(1)), acc))
        return rules

    # ---- run + compare ------------------------------------------------------------------------------------
    def run(self, golden, mode="fast", subtest=(), make_lattice=None, floor=0.0, **replay_options):
        """Replay one subtest and apply the compare rules.  golden: {"<data>-t<time>": array}.  An `accuracy` is an
        absolute bound as with h5diff -d; rules without one demand equal bits (the reference's strict mode), unless
        `floor` > 0 replaces them (a GPU run is "another compiler": the reference itself relaxes such comparisons).
        Returns [(data, time, max abs difference, bound)]; raises AssertionError on the first violation."""
        p = P.parse_file(os.path.join(self.root, self.input_file), self.overrides(mode) + list(subtest))
        if self.checkpoint:  # run.py: `-r <name>`; the restore path is relative to the test's directory
            replay_options = dict(replay_options, restore=self.checkpoint["name"], base_dir=self.root)
        out = R.replay(p, make_lattice=make_lattice, **replay_options)
        sim = P.replace_fname_tokens(p["io.simulationName"], p)
        report = []
        for data, t, acc in self.compare_rules(mode):
            key = f"{data}-t{t}"
            if key not in golden:
                continue  # (a golden the fixtures do not carry, e.g. the t = 10000 dump of the laplace test)
            got, ref = out[f"{data}-{sim}-t{t:08d}"], golden[key]
            if ref.shape != got.shape:  # goldens of runs that are uniform in x and y hold one line along z
                got = got[0, 0]
            bound = max(acc or 0.0, floor)
            diff = float(np.max(np.abs(got - ref)))
            assert diff <= bound, f"{self.name}: {data} at t = {t}: max abs difference {diff:.3e} > {bound:.3e}"
            report.append((data, t, diff, bound))
        if not report:
            raise SpecError(f"{self.name}: no comparison of mode '{mode}' has a golden")
        return report
