"""dqmc_b200.params.Parameters against the reference's own utility::parameters (include/utility.h:50-276,
compiled unmodified into oracle/_ref) on a deliberately awkward file, and on the shipped example's text."""
import os

import pytest

from dqmc_b200.params import Parameters

TRICKY = """
top = 1            ; before any section -> section "global"
[ Lattice ]
L1 = 6   # trailing comment
L2=  10_0
  [hubbard]
U = 4.0e0
t = "1.5"
mu = '-0.1'
name = "quoted # not a comment?"
[simulation]
beta = 4.0abc
nt = 40.9
flag1 = TRUE
flag2 = Off
flag3 = maybe
empty =
quote = "
[ParallelTempering]
betas = 5.0, 4.5,, 4_0.0 ,3.5x, 
bad = 1.0, two
novalue
"""

EXAMPLE = """[Lattice]
L1 = 6
L2 = 6

[hubbard]
U = 4.0                        # On-site interaction strength
t =  1.0                        # Hopping parameters
mu = -0.1                       # Chemical potential

[simulation]
beta = 4.0                      # Inverse temperature
nt = 40                        # Number of Imaginary time step
n_therms = 2000                 # Number of thermalization sweeps
n_sweeps = 40                   # Number of Monte Carlo sweeps
n_bins = 1000                    # Number of total bins. each bin contain average observable after n_sweeps 
n_stab = 10                      # Length of numerical stabilization in imaginary time propagation
symmetric = true
isMeasureUnequalTime = false     # whether to measure unequaltime

[ParallelTempering]
enabled = false
sweep_steps = 20
betas = 5.0, 4.5, 4.0, 3.5, 3.0, 2.5"""

QUERIES = [
    ("global", "top"), ("Lattice", "L1"), ("Lattice", "L2"), ("hubbard", "U"), ("hubbard", "t"), ("hubbard", "mu"),
    ("hubbard", "name"), ("simulation", "beta"), ("simulation", "nt"), ("simulation", "flag1"), ("simulation", "flag2"),
    ("simulation", "flag3"), ("simulation", "empty"), ("simulation", "quote"), ("ParallelTempering", "betas"),
    ("ParallelTempering", "bad"), ("ParallelTempering", "novalue"), ("nosuch", "x"), ("Lattice", "nosuch"),
    ("simulation", "n_bins"), ("simulation", "symmetric"), ("ParallelTempering", "enabled"), ("ParallelTempering", "sweep_steps"),
]


def _mine(p, section, key, kind):
    return [p.getString, p.getInt, p.getDouble, p.getBool, p.getDoubleVector, lambda s, k: p.hasSection(s), p.hasKey][kind](section, key)


@pytest.mark.parametrize("text", [TRICKY, EXAMPLE])
def test_parser_matches_reference(text, tmp_path):
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    path = str(tmp_path / "parameters.in")
    with open(path, "w") as f:
        f.write(text)
    p = Parameters(path)
    for section, key in QUERIES:
        for kind in range(7):
            try:
                want = ref.params_query(path, section, key, kind)
            except RuntimeError:
                want = RuntimeError
            try:
                got = _mine(p, section, key, kind)
            except RuntimeError:
                got = RuntimeError
            assert got == want, (section, key, kind, got, want)


def test_defaults_and_errors(tmp_path):
    p = Parameters(text=EXAMPLE)
    assert p.getInt("simulation", "nt") == 40 and p.getDouble("hubbard", "mu") == -0.1
    assert p.getBool("simulation", "symmetric") is True and p.getBool("ParallelTempering", "enabled", True) is False
    assert p.getDoubleVector("ParallelTempering", "betas") == [5.0, 4.5, 4.0, 3.5, 3.0, 2.5]
    assert p.getInt("simulation", "missing", 7) == 7 and p.getBool("nosuch", "k", True) is True
    assert p.getDouble("simulation", "symmetric", 2.5) == 2.5  # conversion failure -> default, like the catch-all
    with pytest.raises(RuntimeError):
        p.getInt("simulation", "missing")
    with pytest.raises(RuntimeError):
        Parameters(str(tmp_path / "does_not_exist.in"))
    assert p.set("x", "y", 3).getInt("x", "y") == 3
