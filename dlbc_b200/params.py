"""Reader for the reference's parameter files (``*.in``) -- SURVEY.md 8(f) row 1.

Mirrors src/dlbc/parameters.d of the reference:

* ``//`` starts a comment (parameters.d:222-225); ``[section]`` sets the prefix of the keys that follow
  (:241-246); ``key = value`` assigns ``section.key`` (:227-238).
* The qualified name of a parameter is its module path below ``dlbc`` with a repeated last component
  dropped (makeQualModuleName, :544-551): ``dlbc.lb.lb.tau`` is ``lb.tau``, ``dlbc.io.io.outputPath``
  is ``io.outputPath``, ``dlbc.lb.force.gcc`` is ``lb.force.gcc``.
* An unknown key is a fatal error that names the line (:467-479); ``--parameter name=value`` overrides
  are applied after the file with line number 0 ("from command line").
* Values are D literals: ``42``, ``1.0``, ``true``, bare or quoted strings, bare enum members
  (``BDist2``), and (nested) arrays ``[ a, b ]``.

KNOWN lists every ``@("param")`` variable of the reference (name -> (type, default)), so that any of its
parameter files parses; only the lattice-Boltzmann subset is consumed by `dlbc_b200.replay`.
"""
from __future__ import annotations

import re

# enum members (reference declaration order = integer value)
ENUMS = {
    "EqDistForm": ["SecondOrder", "ThirdOrder", "BDist2"],                                   # lb/eqdist.d:23-36
    "PsiForm": ["Linear", "Exponential"],                                                    # lb/force.d:48-57
    "MaskInit": ["None", "File", "Plates", "Tube", "Box", "Walls"],                          # lb/mask.d:51-79
    "Axis": ["X", "Y", "Z"],                                                                  # lb/connectivity.d:107-120
    "FluidInit": ["None", "File", "Const", "EqDist", "EqDistPerturb", "EqDistPerturbFrac", "ConstRandom",
                  "EqDistRandom", "EqDistSphere", "EqDistSphereFrac", "EqDistTwoSpheres",
                  "EqDistTwoSpheresFrac", "EqDistCylinder", "EqDistCylinderFrac", "EqDistLamellae",
                  "EqDistLamellaeFrac"],                                                     # lb/init.d:140-260
    "FileFormat": ["Ascii", "HDF5"],
    "ElecChargeInit": None, "ElecDielInit": None, "PoissonSolver": None, "BoundaryPhi": None,  # any identifier
}

# qualified name -> (type, default).  Types: int double bool string size_t, enum:<Name>, with [] / [][] suffixes.
KNOWN = {
    # dlbc.lattice / dlbc.parallel / dlbc.random / dlbc.timers
    "lattice.gn": ("size_t[]", []),
    "parallel.nc": ("int[]", []), "parallel.showTopology": ("bool", False),
    "random.seed": ("int", 0), "random.shiftSeedByRank": ("bool", True),
    "timers.enableIO": ("bool", False),
    # dlbc.lb.lb
    "lb.timesteps": ("int", 0), "lb.components": ("int", 1), "lb.fieldNames": ("string[]", []),
    "lb.tau": ("double[]", []),
    # dlbc.lb.eqdist / force
    "lb.eqdist.eqDistForm": ("enum:EqDistForm", "SecondOrder"),
    "lb.force.globalAcc": ("double[]", []), "lb.force.enableShanChen": ("bool", False),
    "lb.force.gcc": ("double[][]", []), "lb.force.gwc": ("double[]", []),
    "lb.force.psiForm": ("enum:PsiForm", "Linear"),
    # dlbc.lb.init
    "lb.init.fluidInit": ("enum:FluidInit[]", []), "lb.init.fluidDensities": ("double[][]", []),
    "lb.init.fluidPerturb": ("double[]", []), "lb.init.randomReseed": ("bool[]", []),
    "lb.init.randomSeed": ("int[]", []), "lb.init.initRadius": ("double", 0.0),
    "lb.init.initOffset": ("double[]", []), "lb.init.initSeparation": ("double[]", []),
    "lb.init.initAxis": ("enum:Axis", "X"), "lb.init.interfaceThickness": ("double", 0.0),
    "lb.init.lamellaeWidths": ("double[]", []), "lb.init.fluidFiles": ("string[]", []),
    # dlbc.lb.mask
    "lb.mask.maskInit": ("enum:MaskInit", "None"), "lb.mask.maskFile": ("string", ""),
    "lb.mask.initAxis": ("enum:Axis", "X"), "lb.mask.wallOffset": ("int", 0),
    "lb.mask.wallThickness": ("int", 1), "lb.mask.dimensionHasWall": ("bool[][]", []),
    # dlbc.lb.io / laplace
    "lb.io.populationFreq": ("int", 0), "lb.io.densityFreq": ("int", 0), "lb.io.colourFreq": ("int", 0),
    "lb.io.velocityFreq": ("int", 0), "lb.io.forceFreq": ("int", 0), "lb.io.maskFreq": ("int", 0),
    "lb.io.laplaceFreq": ("int", 0),
    "lb.laplace.startCheck": ("int", -1), "lb.laplace.relAccuracy": ("double", 0.0),
    # dlbc.io.*
    "io.enableIO": ("bool", True), "io.simulationName": ("string", ""), "io.outputFormat": ("enum:FileFormat", "HDF5"),
    "io.outputPath": ("string", "."), "io.createPath": ("bool", True), "io.startOutput": ("int", 0),
    "io.checkpoint.cpPath": ("string", "."), "io.checkpoint.cpRestorePath": ("string", ""),
    "io.checkpoint.cpFreq": ("int", 0), "io.checkpoint.cpKeep": ("int", 1),
    "io.hdf5.writeChunked": ("bool", False),
    # dlbc.elec.* (electrokinetics: parsed so that every reference file loads; not on this path)
    "elec.enableElec": ("bool", False), "elec.localDiel": ("bool", False), "elec.dielGlobal": ("double", 0.0),
    "elec.fluidOnElec": ("bool", False), "elec.elecOnFluid": ("bool", False), "elec.beta": ("double", 1.0),
    "elec.externalField": ("double[]", []), "elec.fluidDiel": ("double[]", []), "elec.solidDiel": ("double", 1.0),
    "elec.boundaryPhi": ("enum:BoundaryPhi[][]", []), "elec.dropPhi": ("double[][]", []),
    "elec.init.chargeInit": ("enum:ElecChargeInit", ""), "elec.init.chargeSolid": ("double", 0.0),
    "elec.init.chargeDensitySolid": ("double", 0.0), "elec.init.saltConc": ("double", 0.0),
    "elec.init.lamellaeWidths": ("double[]", []), "elec.init.chargeLamellae": ("double[]", []),
    "elec.init.dielInit": ("enum:ElecDielInit", ""), "elec.init.dielUniform": ("double", 0.0),
    "elec.init.bjerrumLength": ("double", 0.0), "elec.init.dielCapacitor": ("double", 0.0),
    "elec.init.dielCapacitor2": ("double", 0.0), "elec.init.initAxis": ("enum:Axis", "X"),
    "elec.init.initMaxIterations": ("int", 0), "elec.init.initFluxToleranceRel": ("double", 0.0),
    "elec.force.enableElectrostatic": ("bool", True), "elec.force.enableDielectrophoretic": ("bool", True),
    "elec.flux.thermalDiffusionCoeff": ("double", 0.0), "elec.flux.deltaTDC": ("double", 0.0),
    "elec.flux.deltaMuPos": ("double", 0.0), "elec.flux.deltaMuNeg": ("double", 0.0),
    "elec.flux.enableDiffusiveFlux": ("bool", True), "elec.flux.enableAdvectiveFlux": ("bool", True),
    "elec.io.chargeFreq": ("int", 0), "elec.io.potFreq": ("int", 0), "elec.io.dielFreq": ("int", 0),
    "elec.io.fieldFreq": ("int", 0),
    "elec.poisson.solver": ("enum:PoissonSolver", ""), "elec.poisson.sorMaxIterations": ("int", 0),
    "elec.poisson.sorCheckIterations": ("int", 0), "elec.poisson.sorShowIterations": ("int", 0),
    "elec.poisson.sorToleranceRel": ("double", 0.0),
}


class ParameterError(ValueError):
    """What the reference reports with writeLogF (fatal)."""


# ---- D literal parsing ---------------------------------------------------------------------------
_TOKEN = re.compile(r'\s*(\[|\]|,|"(?:[^"\\]|\\.)*"|[^\[\],\s"]+)')


def _tokens(text):
    pos, out = 0, []
    text = text.strip()
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            raise ParameterError(f"cannot tokenise {text[pos:]!r}")
        out.append(m.group(1))
        pos = m.end()
    return out


def _tree(tokens):
    """tokens -> nested lists of scalar strings"""
    def node(i):
        if tokens[i] == "[":
            items, i = [], i + 1
            while tokens[i] != "]":
                item, i = node(i)
                items.append(item)
                if tokens[i] == ",":
                    i += 1
            return items, i + 1
        if tokens[i] in ("]", ","):
            raise ParameterError(f"unexpected {tokens[i]!r}")
        return tokens[i], i + 1
    try:
        val, i = node(0)
    except IndexError:
        raise ParameterError("unbalanced brackets") from None
    if i != len(tokens):
        raise ParameterError(f"trailing input {tokens[i:]!r}")
    return val


def _scalar(tok, base):
    if base in ("int", "size_t"):
        return int(tok)
    if base == "double":
        return float(tok)
    if base == "bool":
        if tok.lower() not in ("true", "false"):  # std.conv parses bool case-insensitively ("True" in init-2d/*.in)
            raise ParameterError(f"{tok!r} is not a bool (true/false)")
        return tok.lower() == "true"
    if base == "string":
        return tok[1:-1] if len(tok) >= 2 and tok[0] == '"' and tok[-1] == '"' else tok
    if base.startswith("enum:"):
        members = ENUMS[base[5:]]
        if members is not None and tok not in members:
            raise ParameterError(f"{tok!r} is not a member of {base[5:]} {members}")
        return tok
    raise ParameterError(f"unknown parameter type {base}")


def convert(value_string, type_):
    """std.conv.to!T of a D literal."""
    depth = type_.count("[]")
    base = type_.replace("[]", "")
    if depth == 0:
        toks = _tokens(value_string)
        if len(toks) != 1 and base != "string":
            raise ParameterError(f"expected one {base}, got {value_string!r}")
        return _scalar(toks[0] if base != "string" else value_string.strip(), base)
    tree = _tree(_tokens(value_string))

    def walk(t, d):
        if d == 0:
            if isinstance(t, list):
                raise ParameterError(f"array too deep for {type_}")
            return _scalar(t, base)
        if not isinstance(t, list):
            raise ParameterError(f"expected an array for {type_}, got {t!r}")
        return [walk(x, d - 1) for x in t]
    return walk(tree, depth)


# ---- files ---------------------------------------------------------------------------------------
def defaults():
    return {k: (list(v[1]) if isinstance(v[1], list) else v[1]) for k, v in KNOWN.items()}


def _assign(params, key, value_string, ln):
    if key not in KNOWN:
        if ln > 0:
            raise ParameterError(f"Unknown key at line {ln}: '{key}'.")     # parameters.d:472
        raise ParameterError(f"Unknown key from command line: '{key}'.")     # parameters.d:475
    try:
        params[key] = convert(value_string, KNOWN[key][0])
    except (ParameterError, ValueError) as e:
        where = f"line {ln}" if ln > 0 else "command line"
        raise ParameterError(f"{where}: cannot convert {value_string!r} for '{key}' ({KNOWN[key][0]}): {e}") from None


def parse_text(text, overrides=(), params=None):
    """Parse a parameter file's text, then apply ``name=value`` overrides (--parameter).  Returns a dict
    qualified name -> value holding every known parameter (defaults where unset) plus ``_set``: the
    names that were assigned explicitly."""
    params = defaults() if params is None else params
    given = params.setdefault("_set", set())
    section = ""
    for ln, raw in enumerate(text.splitlines(), 1):
        line = raw.split("//", 1)[0]
        eq = line.find("=")
        if eq > 0:
            key, val = line[:eq].strip(), line[eq + 1:].strip()
            key = f"{section}.{key}" if section else key
            _assign(params, key, val, ln)
            given.add(key)
        else:
            for m in re.finditer(r"\[(.*?)\]", line):
                section = m.group(1)
    for ov in overrides:
        if "=" not in ov:
            raise ParameterError(f"--parameter expects name=value, got {ov!r}")
        key, val = ov.split("=", 1)
        _assign(params, key.strip(), val.strip(), 0)
        given.add(key.strip())
    return params


def read_lines(path):
    """The lines of a parameter file.  An HDF5 file written by the reference (or by dlbc_b200.h5lite.write_file)
    carries the parameter file it was produced with as the `input` attribute of its root group and can be used
    in its place (parameters.d:172-174,339-351; io/hdf5.d:529-548)."""
    if str(path).endswith(".h5"):
        from . import h5lite
        attrs = h5lite.read_attributes(path)
        if "input" not in attrs:
            raise ParameterError(f"{path}: no 'input' attribute to read the parameters from")
        return [str(x) for x in attrs["input"]]
    with open(path) as f:
        return f.read().splitlines()


def parse_file(path, overrides=()):
    return parse_text("\n".join(read_lines(path)) + "\n", overrides)


def _to_string(v):
    """std.conv.to!string of a scalar parameter value (doubles print as %g)."""
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, float):
        return "%g" % v
    return str(v)


def replace_fname_tokens(name, params):
    """replaceFnameTokens (parameters.d:507-527): ``%lb.force.gcc[1][0]%`` in a file or simulation name is
    replaced by the value of that parameter (or array element)."""
    def sub(m):
        key, idx = m.group(1), [int(i) for i in re.findall(r"\[(\d+)\]", m.group(2))]
        if key not in KNOWN:
            return m.group(0)
        v = params[key]
        try:
            for i in idx:
                v = v[i]
        except (IndexError, TypeError):
            raise ParameterError(f"token {m.group(0)} does not address an element of '{key}'") from None
        return _to_string(v)
    return re.sub(r"%([A-Za-z][A-Za-z0-9_.]*)((?:\[\d+\])*)%", sub, name)
