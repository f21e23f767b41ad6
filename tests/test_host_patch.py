"""pencil-code_b200/fortran/host.patch (the lines of the reference host the Fortran wrapper needs) applies
cleanly to the reference's sources.  Needs /root/reference (this container only) and `patch`; skipped elsewhere."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
PATCH = os.path.join(ROOT, "pencil-code_b200", "fortran", "host.patch")
FILES = ["entropy.f90", "magnetic.f90", "viscosity.f90", "hydro.f90", "equ.f90", "timestep.f90", "noentropy.f90", "nomagnetic.f90",
         "noforcing.f90"]


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")) or shutil.which("patch") is None, reason="reference tree or patch(1) not here")
def test_host_patch_applies_to_the_reference(tmp_path):
    os.makedirs(tmp_path / "src")
    for f in FILES:
        shutil.copy(os.path.join(REF, "src", f), tmp_path / "src" / f)
    r = subprocess.run(["patch", "-p1", "--dry-run", "-i", PATCH], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    touched = [ln.split()[-1] for ln in r.stdout.splitlines() if ln.startswith("checking file")]
    assert sorted(os.path.basename(t) for t in touched) == sorted(FILES)
    assert "FAILED" not in r.stdout and "fuzz" not in r.stdout


def test_wrapper_only_imports_what_the_reference_exports():
    """gpu_b200.f90 `use`s module variables only where the reference's *.h public lists export them; private run parameters
    come through pushpars2c"""
    src = open(os.path.join(ROOT, "pencil-code_b200", "fortran", "gpu_b200.f90")).read().lower()
    for bad in ("use viscosity, only: nu", "use magnetic, only: eta", "use timestep, only: alpha_ts", "use hydro, only: omega",
                "use energy, only: lupw_ss"):
        assert bad not in src, bad
    assert "viscosity_pars => pushpars2c" in src and "energy_pars => pushpars2c" in src and "magnetic_pars => pushpars2c" in src
