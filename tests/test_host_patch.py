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
         "noforcing.f90", "particles_dust.f90", "noparticles.f90", "particles_main.f90", "noparticles_main.f90"]


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


# module of gpu_b200.f90's `use` statements -> the source files that can fill that slot in a build the wrapper supports
MODULE_FILES = {
    "equationofstate": ["eos_idealgas.f90"], "hydro": ["hydro.f90"], "density": ["density.f90"],
    "viscosity": ["viscosity.f90"], "magnetic": ["magnetic.f90", "nomagnetic.f90"], "energy": ["entropy.f90", "noentropy.f90"],
    "gravity": ["gravity_simple.f90", "nogravity.f90"], "forcing": ["forcing.f90", "noforcing.f90"], "mpicomm": ["mpicomm.f90", "nompicomm.f90"],
    "particles": ["particles_dust.f90", "noparticles.f90"], "particles_main": ["particles_main.f90", "noparticles_main.f90"],
    "particles_cdata": ["particles_cdata.f90"], "general": ["general.f90"], "messages": ["messages.f90"],
}


def _public_names(path):
    """names a module file makes public: None when the module has no `private` default (everything is public)"""
    import re
    src = open(path, errors="replace").read()
    head = src.split("\n  contains")[0] if "\n  contains" in src else src.split("\ncontains")[0]
    text = head
    for inc in re.findall(r"include\s+'([^']+)'", head):
        ip = os.path.join(os.path.dirname(path), inc)
        if os.path.exists(ip):
            text += "\n" + open(ip, errors="replace").read()
    text = re.sub(r"!.*", "", text)
    text = re.sub(r"&\s*\n\s*&?", " ", text)
    if not re.search(r"^\s*private\s*$", text, re.M | re.I):
        return None
    names = set()
    for m in re.finditer(r"^\s*public\s*::\s*(.*)$", text, re.M | re.I):
        names.update(n.strip().lower() for n in m.group(1).split(",") if n.strip())
    for m in re.finditer(r"^[^\n]*,\s*public\b[^\n]*::\s*(.*)$", text, re.M | re.I):           # attribute form
        names.update(re.match(r"\w+", n.strip()).group(0).lower() for n in m.group(1).split(",") if n.strip())
    return names


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")) or shutil.which("patch") is None, reason="reference tree or patch(1) not here")
def test_every_name_the_wrapper_imports_is_public_in_the_patched_host(tmp_path):
    """every `use M, only: ...` of gpu_b200.f90 against the public list of every source file that can fill slot M
    (the module's own `public ::` lines + its include'd *.h), after host.patch"""
    import re
    os.makedirs(tmp_path / "src")
    for f in os.listdir(os.path.join(REF, "src")):
        if f.endswith((".f90", ".h")):
            shutil.copy(os.path.join(REF, "src", f), tmp_path / "src" / f)
    for f in os.listdir(tmp_path / "src"):
        os.chmod(tmp_path / "src" / f, 0o644)
    r = subprocess.run(["patch", "-p1", "-i", PATCH], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    src = open(os.path.join(ROOT, "pencil-code_b200", "fortran", "gpu_b200.f90")).read()
    src = re.sub(r"!.*", "", src)
    src = re.sub(r"&\s*\n\s*&?", " ", src)
    uses = re.findall(r"^\s*use\s+(\w+)\s*,\s*only\s*:\s*(.*)$", src, re.M | re.I)
    assert len(uses) >= 12
    missing = []
    for mod, names in uses:
        mod = mod.lower()
        if mod == "cdata":
            continue
        assert mod in MODULE_FILES, mod
        for item in names.split(","):
            name = item.split("=>")[-1].strip().lower()                  # local => remote
            for f in MODULE_FILES[mod]:
                pub = _public_names(str(tmp_path / "src" / f))
                if pub is not None and name not in pub:
                    missing.append((mod, name, f))
    assert not missing, missing
    # whole-module imports only of modules without a private default
    for mod in re.findall(r"^\s*use\s+(\w+)\s*$", src, re.M | re.I):
        if mod.lower() in ("cdata",):
            continue
        assert mod.lower() in MODULE_FILES, mod
        for f in MODULE_FILES[mod.lower()]:
            assert _public_names(str(tmp_path / "src" / f)) is None, (mod, f)


def test_fortran_mirrors_of_the_abi_structs_match_the_header():
    """type, bind(C) :: pcb200_params / pcb200_particle_params / pcb200_forcing of gpu_b200.f90 against include/pcb200.h:
    same members in the same order, integer vs real, array extents (reversed: Fortran is column-major)"""
    import re
    h = open(os.path.join(ROOT, "include", "pcb200.h")).read()
    f = open(os.path.join(ROOT, "pencil-code_b200", "fortran", "gpu_b200.f90")).read()
    consts = {k: v for k, v in re.findall(r"#define\s+(PCB200_\w+)\s+(\d+)", h)}

    def c_members(name):
        body = re.search(r"typedef struct " + name + r" \{(.*?)\} " + name + ";", h, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        out = []
        for decl in body.split(";"):
            if not decl.strip():
                continue
            ctype, rest = decl.strip().split(None, 1)
            for v in rest.split(","):
                m = re.match(r"\s*(\w+)((?:\[[^\]]+\])*)", v)
                dims = [consts.get(d.strip(), d.strip()) for d in re.findall(r"\[([^\]]+)\]", m.group(2))]
                out.append(("integer" if "int" in ctype else "real", m.group(1).lower(), dims))
        return out

    def f_members(name):
        body = re.search(r"type, bind\(C\) :: " + name + r"\n(.*?)endtype", f, re.S).group(1)
        fconst = {k.upper(): v for k, v in re.findall(r"(PCB200_\w+)=(\d+)", f)}
        out = []
        for ln in body.splitlines():
            ln = ln.split("!")[0].strip()
            if not ln:
                continue
            ftype, rest = ln.split("::")
            for n, d in re.findall(r"(\w+)(\([^)]*\))?", rest):
                dims = [fconst.get(x.strip().upper(), x.strip()) for x in d.strip("()").split(",")] if d else []
                out.append(("integer" if ftype.strip().startswith("integer") else "real", n.lower(), dims[::-1]))
        return out

    for name in ("pcb200_params", "pcb200_particle_params", "pcb200_forcing"):
        c, fm = c_members(name), f_members(name)
        assert len(c) == len(fm) and len(c) > 3, (name, len(c), len(fm))
        for a, b in zip(c, fm):
            assert a == b, (name, a, b)


def test_wrapper_block_structure_is_balanced():
    """no Fortran compiler in this image: at least every block of gpu_b200.f90 that opens closes with its own keyword, in
    order (module/type/interface/subroutine/function/if-then/do/select/where), continuation lines joined"""
    import re
    src = open(os.path.join(ROOT, "pencil-code_b200", "fortran", "gpu_b200.f90")).read()
    lines, cur = [], ""
    for ln in src.splitlines():
        ln = re.sub(r"'[^']*'", "''", ln)
        ln = ln.split("!")[0].rstrip()
        if not ln.strip():
            continue
        if ln.rstrip().endswith("&"):
            cur += ln.rstrip()[:-1] + " "
            continue
        lines.append((cur + ln).strip().lower())
        cur = ""
    stack = []
    for full in lines:
        for ln in (s.strip() for s in full.split(";")):            # `do ...; ...; enddo` on one line
            if not ln:
                continue
            m = re.match(r"end\s*(module|type|interface|subroutine|function|if|do|select|where)\b", ln)
            if m:
                assert stack and stack[-1] == m.group(1), (ln, stack[-3:])
                stack.pop()
            elif re.match(r"module\s+\w+$", ln):
                stack.append("module")
            elif re.match(r"type\s*(,\s*bind\(c\))?\s*::", ln):
                stack.append("type")
            elif ln == "interface":
                stack.append("interface")
            elif re.match(r"(\w+\s+)*subroutine\s+\w+", ln):
                stack.append("subroutine")
            elif re.match(r"([\w()=]+\s+)*function\s+\w+\s*\(", ln):
                stack.append("function")
            elif re.match(r"(\w+\s*:\s*)?if\s*\(.*\)\s*then$", ln):
                stack.append("if")
            elif re.match(r"(\w+\s*:\s*)?do(\s+\w+\s*=.*)?$", ln):
                stack.append("do")
            elif re.match(r"select\s+case", ln):
                stack.append("select")
            elif re.match(r"where\s*\([^)]*\)$", ln):
                stack.append("where")
    assert stack == [], stack


def test_wrapper_interface_block_matches_the_c_prototypes():
    """every bind(C) interface of gpu_b200.f90 against its prototype in include/pcb200.h: same number of arguments,
    by-value vs by-reference, and int / double / char / struct per argument"""
    import re
    h = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "pcb200.h")).read(), flags=re.S)
    protos = {}
    for ret, name, args in re.findall(r"^\s*(const char \*|int|void)\s*(pcb200_\w+)\s*\(([^)]*)\)\s*;", h, re.M):
        out = []
        for a in [x.strip() for x in args.split(",") if x.strip() and x.strip() != "void"]:
            byref = "*" in a or "[" in a
            base = "char" if re.search(r"\bchar\b", a) else "double" if "double" in a else "float" if "float" in a else \
                   "struct" if re.search(r"pcb200_\w+", a) else "int"
            out.append((base, byref))
        protos[name] = (ret.strip(), out)
    f = open(os.path.join(ROOT, "pencil-code_b200", "fortran", "gpu_b200.f90")).read()
    block = re.search(r"\n  interface\n(.*?)\n  endinterface", f, re.S).group(1)
    block = re.sub(r"&\s*\n\s*&?", " ", block)
    n = 0
    for m in re.finditer(r"function\s+(\w+)\s*\(([^)]*)\)\s*bind\(C,name='(\w+)'\)[^\n]*\n(.*?)endfunction", block, re.S | re.I):
        fname, fargs, cname, body = m.group(1), [a.strip() for a in m.group(2).split(",") if a.strip()], m.group(3), m.group(4)
        assert fname == cname and cname in protos, cname
        decl = {}
        for ln in body.splitlines():
            if "::" not in ln or ln.strip().startswith("import"):
                continue
            spec, names = ln.split("::")
            spec = spec.lower()
            base = "char" if "character" in spec else "double" if "c_double" in spec else "float" if "c_float" in spec else \
                   "struct" if "type(pcb200" in spec else "int"
            for nm in names.split(","):
                decl[nm.strip()] = (base, "value" not in spec)
        got = [decl[a] for a in fargs]
        assert got == protos[cname][1], (cname, got, protos[cname][1])
        n += 1
    assert n >= 15


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not here")
def test_every_free_identifier_of_the_wrapper_exists_in_the_host():
    """names gpu_b200.f90 takes from whole-module imports (Cdata incl. Cparam and the generated cparam.inc switches,
    Particles_cdata): each must be declared in cdata.f90 / cparam.f90 / particles_cdata.f90 or be a module switch that
    some src/*.f90 header announces with `! CPARAM logical, parameter :: l... =`"""
    import glob
    import re
    src = open(os.path.join(ROOT, "pencil-code_b200", "fortran", "gpu_b200.f90")).read()
    code = re.sub(r"!.*", "", src)
    code = re.sub(r"'[^']*'", "''", code)
    code = re.sub(r"&\s*\n\s*&?", " ", code)
    low = code.lower()
    ids = set(re.findall(r"[a-z_]\w*", low))
    declared = set()
    for m in re.finditer(r"::\s*([^\n]*)", low):                              # local declarations and struct members
        declared.update(re.findall(r"[a-z_]\w*", m.group(1)))
    for m in re.finditer(r"(?<![a-z])(?:subroutine|function|module|type(?:,\s*bind\(c\))?\s*::)[ \t]+(\w+)", low):
        declared.add(m.group(1))
    for m in re.finditer(r"(?:subroutine|function)\s+\w+\s*\(([^)]*)\)", low):
        declared.update(a.strip() for a in m.group(1).split(","))
    for m in re.finditer(r"use\s+(\w+)\s*,\s*only\s*:\s*([^\n]*)", low):      # explicit imports (checked by the test above)
        declared.add(m.group(1))
        declared.update(re.findall(r"[a-z_]\w*", m.group(2)))
    declared.update(m.group(1) for m in re.finditer(r"use\s*(?:,\s*intrinsic\s*::)?\s*(\w+)", low))
    host = ""
    for f in ("src/cdata.f90", "src/cparam.f90", "src/particles_cdata.f90", "samples/conv-slab/src/cparam.local"):   # cparam.f90 includes cparam.local
        host += open(os.path.join(REF, f), errors="replace").read().lower()
    host_ids = set(re.findall(r"[a-z_]\w*", host))
    switches = set()
    for f in glob.glob(os.path.join(REF, "src", "*.f90")):
        head = open(f, errors="replace").read(6000).lower()
        switches.update(re.findall(r"!\s*cparam\s+logical\s*,\s*parameter\s*::\s*(l\w+)", head))
    keywords = set("""use only intrinsic implicit none include integer parameter type bind c endtype real logical character len
        kind dimension save pointer null contains subroutine endsubroutine function endfunction result import value intent in out
        inout interface endinterface if then else endif do enddo exit call return select case default endselect trim merge int all
        mod exp transfer associated nullify where elsewhere endwhere target contiguous module endmodule and or not true false name
        c_int c_int32_t c_double c_char c_ptr c_null_char c_null_ptr c_f_pointer _c_int""".split())
    free = sorted(i for i in ids if i not in declared and i not in keywords and i not in host_ids and i not in switches)
    assert free == [], free
