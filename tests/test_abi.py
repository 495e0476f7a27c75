"""CPU tests of the drop-in boundary: the C-ABI library loads without a GPU, exports every
symbol include/dlesm_b200.h declares (and nothing undeclared under the dlesm_ prefix), the
Python binding covers all of them, and device entry points fail loudly -- never fall back."""
import ctypes as C
import os
import re
import subprocess

import pytest

import dl_esm_inf_b200 as fd
from dl_esm_inf_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dlesm_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(dlesm_[a-z0-9_]+)\s*\(", text))


def exported_symbols():
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True,
                         check=True).stdout
    return {ln.split()[-1] for ln in out.splitlines() if " T " in ln and ln.split()[-1].startswith("dlesm_")}


def test_header_library_and_binding_agree():
    decl, exp, bound = declared_symbols(), exported_symbols(), set(_lib.SIGNATURES)
    assert decl == exp, (sorted(decl - exp), sorted(exp - decl))
    assert decl == bound, (sorted(decl - bound), sorted(bound - decl))
    L = _lib.lib()
    for name in decl:
        assert hasattr(L, name)


def test_no_torch_or_cxx_types_in_the_abi():
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)   # code only
    assert "torch" not in text and "std::" not in text and "template" not in text
    assert 'extern "C"' in text


def test_product_never_touches_the_oracle():
    """the oracle is test infrastructure: nothing under dl_esm_inf_b200/ may import, link or
    load it (a product path through the oracle would void every parity claim)"""
    pkg = os.path.join(ROOT, "dl_esm_inf_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".f90", ".F90")):
                src = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "liboracle" not in src and "from oracle" not in src and "import oracle" not in src, fn
    ldd = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in ldd


def test_device_entry_points_fail_loudly_without_gpu():
    L = _lib.lib()
    p = C.c_void_p()
    rc = L.dlesm_dev_field_alloc(8, 8, C.byref(p))
    if rc == 0:   # a GPU is present (running on the B200 box): nothing to check here
        L.dlesm_dev_field_free(p)
        pytest.skip("GPU present")
    assert rc == -20 and b"no CPU fallback" in L.dlesm_last_error()
    fd.gocean_initialise(1, 1)
    g = fd.grid_type(fd.GO_ARAKAWA_C, (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE), fd.GO_OFFSET_NE)
    g.decompose(4, 10)
    fd.grid_init(g, 1.0, 1.0)
    with pytest.raises(fd.GoceanStop) as e:
        fd.r2d_field(g, fd.GO_T_POINTS)
    assert e.value.code == -20
    # the sweeps and the deep halo fill refuse pointers they did not allocate: nothing computes on the host
    bogus = [C.c_void_p(0x1000 + 64 * k) for k in range(12)]
    rc = L.dlesm_sw_step2_fused(*bogus, 1.0, 1.0, 180.0, 0.001, 3, 10, 3, 10, None, 0)
    assert rc < 0 and L.dlesm_last_error()
    rc = L.dlesm_sw_step_fused(*bogus[:9], 1.0, 1.0, 180.0, 0.001, 0, 2, 9, 2, 9, None, 0)
    assert rc < 0
    rc = L.dlesm_dev_periodic_halos_deep(bogus[0], 3, 10, 3, 10, 2)
    assert rc < 0


def test_grid_constructor_validation():
    """src/grid_mod.f90:241-278"""
    fd.gocean_initialise(1, 1)
    bcs = (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE)
    with pytest.raises(fd.GoceanStop) as e:
        fd.grid_type(fd.GO_ARAKAWA_C, bcs)                    # offset is "optional" but required
    assert e.value.code == -8
    with pytest.raises(fd.GoceanStop) as e:
        fd.grid_type(7, bcs, fd.GO_OFFSET_NE)
    assert e.value.code == -7
    with pytest.raises(fd.GoceanStop) as e:
        fd.grid_type(fd.GO_ARAKAWA_C, bcs, fd.GO_OFFSET_ANY)
    assert e.value.code == -8
    g = fd.grid_type(fd.GO_ARAKAWA_B, bcs, fd.GO_OFFSET_SE)
    assert g.name == fd.GO_ARAKAWA_B and g.offset == fd.GO_OFFSET_SE


def test_periodic_with_several_ranks_is_refused_like_the_reference():
    """src/grid_mod.f90:437-442, 559-564; allowed only with the explicit extension flag"""
    bcs = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
    fd.gocean_initialise(1, 4)
    fd.set_extensions(0)
    g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, fd.GO_OFFSET_SW)
    g.decompose(16, 16)
    with pytest.raises(fd.GoceanStop) as e:
        fd.grid_init(g, 1.0, 1.0)
    assert e.value.code == -6
    fd.set_extensions(fd.EXT_PERIODIC_DISTMEM)
    try:
        fd.grid_init(g, 1.0, 1.0)
        c = g.comms
        assert c.nsend == 8 and c.nrecv == 8      # wrap-around: every tile has 8 neighbours
    finally:
        fd.set_extensions(0)
        fd.gocean_initialise(1, 1)


def test_fortran_bindings_name_only_exported_symbols():
    """every bind(C, name=...) of fortran/*.f90 is an entry point the shared library exports (no Fortran
    compiler here: this is the link-time check a build would make)"""
    import glob
    import re
    L = _lib.lib()
    names = set()
    for path in glob.glob(os.path.join(ROOT, "fortran", "*.f90")):
        names |= set(re.findall(r'name="(dlesm_\w+)"', open(path).read()))
    assert len(names) >= 40
    missing = [n for n in sorted(names) if not hasattr(L, n)]
    assert not missing, missing


def _fortran_block_errors(txt):
    """the structural check a Fortran front end makes first: every module / interface / type / subroutine /
    function / do / if-then / select / associate is closed by the matching END, in order (free form, comments
    and continuation lines removed).  Calibrated on the reference's 17 source files: no false positive."""
    import re
    lines = ["" if l.lstrip().startswith("!") else l.split("!")[0].rstrip() for l in txt.split("\n")]
    joined = re.sub(r"&\s*\n\s*&?", " ", "\n".join(lines))
    stack, errs = [], []
    for ln, l in enumerate(joined.split("\n"), 1):
        t = l.strip().lower()
        if not t:
            continue
        t = re.sub(r"^\d+\s+", "", t)
        t = re.sub(r"^\w+\s*:\s*(?=(do|if|select)\b)", "", t)
        m = re.match(r"end\s*(module|subroutine|function|interface|type|do|if|select|program|associate|block)\b", t)
        if m:
            if not stack:
                errs.append((ln, "end %s with nothing open" % m.group(1)))
            else:
                top = stack.pop()
                if top != m.group(1):
                    errs.append((ln, "end %s closes %s" % (m.group(1), top)))
            continue
        if re.match(r"end\s*$", t):
            if stack:
                stack.pop()
            continue
        if re.match(r"module\s+(?!procedure)\w+", t):
            stack.append("module")
        elif re.match(r"program\s+\w+", t):
            stack.append("program")
        elif re.match(r"(abstract\s+)?interface\b", t):
            stack.append("interface")
        elif re.match(r"type\s*(,[^:]*)?(::)?\s*\w+\s*$", t) and not re.match(r"type\s*\(", t):
            stack.append("type")
        elif re.match(r"((pure|elemental|recursive|impure)\s+)*subroutine\s+\w+", t):
            stack.append("subroutine")
        elif re.match(r"((pure|elemental|recursive|impure)\s+)*([\w\(\)=\*, ]+\s+)?function\s+\w+\s*\(", t):
            stack.append("function")
        elif re.match(r"do\b(\s+\w+\s*=|\s+while|\s*$|\s+concurrent)", t):
            stack.append("do")
        elif re.match(r"if\s*\(.*\)\s*then\s*$", t):
            stack.append("if")
        elif re.match(r"select\s+(case|type)\b", t):
            stack.append("select")
        elif re.match(r"associate\s*\(", t):
            stack.append("associate")
    errs += [(0, "unclosed %s" % k) for k in stack]
    return errs


def test_fortran_sources_are_structurally_balanced():
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "fortran", "*.f90"))):
        assert _fortran_block_errors(open(path).read()) == [], path
    # the checker itself: an unclosed block and a crossed END are seen
    assert _fortran_block_errors("module m\ncontains\nsubroutine s()\nif (a) then\nend subroutine s\nend module m\n")
    assert _fortran_block_errors("subroutine s()\ndo i = 1, 2\nend subroutine\n")


def test_field_mod_and_grid_mod_patches_apply_to_the_reference(tmp_path):
    """fortran/patches/*.patch turn the reference's field_mod / grid_mod into the B200 build (public
    generics dispatching on data_on_device).  Where the reference tree is present (the build container)
    they must apply cleanly; every dlesm_* name they call must be declared in dlesm_b200_mod.f90."""
    import re
    import shutil
    import subprocess
    mod = open(os.path.join(ROOT, "fortran", "dlesm_b200_mod.f90")).read()
    declared = set(re.findall(r'name="(dlesm_\w+)"', mod)) | {"dlesm_comms"}
    # ... and the module procedures of its `contains` part (dlesm_check, the two hook wrappers)
    declared |= set(re.findall(r"^\s*(?:subroutine|function)\s+(dlesm_\w+)", mod[mod.index("\ncontains"):], flags=re.M | re.I))
    assert {"dlesm_check", "dlesm_read_from_device_hook", "dlesm_write_to_device_hook"} <= declared
    for name in ("field_mod", "grid_mod"):
        patch = os.path.join(ROOT, "fortran", "patches", name + ".f90.patch")
        added = [l[1:] for l in open(patch) if l.startswith("+") and not l.startswith("+++")]
        called = set(re.findall(r"\b(dlesm_\w+)", " ".join(a.split("!")[0] for a in added)))
        assert called and called <= declared | {"dlesm_b200_mod"}, called - declared
    ref = "/root/reference/finite_difference/src"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present on this box")
    work = tmp_path / "finite_difference" / "src"
    work.mkdir(parents=True)
    for name in ("field_mod", "grid_mod"):
        shutil.copy(os.path.join(ref, name + ".f90"), work / (name + ".f90"))
        patch = os.path.join(ROOT, "fortran", "patches", name + ".f90.patch")
        p = subprocess.run(["patch", "-p1", "-d", str(tmp_path), "-i", patch], capture_output=True, text=True)
        assert p.returncode == 0, p.stdout + p.stderr
        out = (work / (name + ".f90")).read_text()
        assert "use dlesm_b200_mod" in out
        assert _fortran_block_errors(out) == [], name      # the patched module still closes every block
    fm = (work / "field_mod.f90").read_text()
    # fld%halo_exchange no longer reaches exchange_generic (-> post_receive, which the NCCL variant of
    # parallel_utils_mod does not have)
    body = fm[fm.index("subroutine halo_exchange(self, depth)"):fm.index("end subroutine halo_exchange")]
    assert "call exchange_generic" not in body and "dlesm_dev_halo_exchange" in body and "dlesm_host_halo_exchange" in body


def _c_prototypes():
    import re
    hdr = open(os.path.join(ROOT, "include", "dlesm_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    hdr = re.sub(r"//[^\n]*", "", hdr)
    protos = {}
    for m in re.finditer(r"\b([A-Za-z_][\w \*]*?)\b(dlesm_\w+)\s*\(([^;{}]*?)\)\s*;", hdr):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if "typedef" in ret:
            continue
        protos[name] = (ret, [] if args in ("void", "") else [a.strip() for a in args.split(",")])
    return protos, hdr


def _c_kind(a):
    """(base type, levels of indirection) of a C parameter declaration"""
    import re
    a = re.sub(r"\b(const|restrict|__restrict__)\b", "", a).strip()
    arr = "[" in a
    a = re.sub(r"\[[^\]]*\]", "", a)
    stars = a.count("*") + (1 if arr else 0)
    toks = a.replace("*", " ").split()
    return (" ".join(toks[:-1]) if len(toks) > 1 else toks[0]), stars


def test_fortran_interfaces_match_the_c_prototypes():
    """the check a Fortran compiler + linker cannot make and this image cannot run at all: every
    bind(C) interface of fortran/*.f90 against the prototype of include/dlesm_b200.h -- the same number of
    arguments, by value where C takes a scalar and by reference where it takes a pointer, c_int / c_double /
    c_long_long / c_size_t / c_ptr against int / double / long long / size_t / pointers, and the result type"""
    import glob
    import re
    protos, _ = _c_prototypes()
    scal = {"integer(c_int)": "int", "integer(c_long_long)": "long long", "real(c_double)": "double",
            "integer(c_size_t)": "size_t", "logical(c_bool)": "bool"}
    bad, n = [], 0
    for path in sorted(glob.glob(os.path.join(ROOT, "fortran", "*.f90"))):
        txt = re.sub(r"&\s*\n\s*&?", " ", open(path).read())
        pat = (r'(function|subroutine)\s+(\w+)\s*\(([^)]*)\)\s*bind\(C,\s*name="(\w+)"\)'
               r'(?:\s*result\((\w+)\))?(.*?)end\s+(?:function|subroutine)')
        for kind, _fname, args, cname, res, body in re.findall(pat, txt, flags=re.S | re.I):
            n += 1
            args = [a.strip() for a in args.split(",") if a.strip()]
            decl = {}
            for line in body.split("\n"):
                line = line.split("!")[0].strip()
                if "::" not in line:
                    continue
                lhs, rhs = line.split("::", 1)
                for v in re.split(r",(?![^()]*\))", rhs):
                    vn = re.match(r"\s*(\w+)", v)
                    if vn:
                        decl[vn.group(1).lower()] = lhs.strip().lower()
            if cname not in protos:
                bad.append((cname, "no prototype in the header"))
                continue
            ret, cargs = protos[cname]
            if len(cargs) != len(args):
                bad.append((cname, "argument count: Fortran %d, C %d" % (len(args), len(cargs))))
                continue
            for a, ca in zip(args, cargs):
                lhs = decl.get(a.lower())
                if lhs is None:
                    bad.append((cname, "dummy argument %s has no declaration" % a))
                    continue
                base, stars = _c_kind(ca)
                byval = "value" in lhs
                head = lhs.split(",")[0].strip()
                if head in scal:
                    ok = base == scal[head] and stars == (0 if byval else 1)
                elif head == "type(c_ptr)":
                    ok = stars == (1 if byval else 2) or (byval and "(*" in ca)
                elif head == "type(c_funptr)":
                    ok = byval
                elif head.startswith("character"):
                    ok = stars == 1 and base in ("char", "void")
                elif head.startswith("type(dlesm_"):
                    ok = stars == 1 and base == head[5:-1]
                else:
                    ok = False
                if not ok:
                    bad.append((cname, "argument %s: Fortran '%s' against C '%s'" % (a, lhs, ca)))
            if kind.lower() == "subroutine":
                if ret != "void":
                    bad.append((cname, "a subroutine, but C returns " + ret))
            else:
                rl = decl.get((res or "").lower(), "").split(",")[0].strip()
                rb, rs = _c_kind(ret + " x")
                if rl in scal and not (rb == scal[rl] and rs == 0):
                    bad.append((cname, "result '%s' against C '%s'" % (rl, ret)))
                if rl == "type(c_ptr)" and rs < 1:
                    bad.append((cname, "result c_ptr against C '%s'" % ret))
    assert n >= 40
    assert not bad, bad


def test_fortran_dlesm_comms_mirrors_the_c_struct():
    """member order of the interoperable derived type == the C struct's == the ctypes mirror's"""
    import re
    _, hdr = _c_prototypes()
    body = re.search(r"typedef struct(?:\s+\w+)?\s*\{([^{}]*?)\}\s*dlesm_comms;", hdr, flags=re.S)
    assert body is not None
    cnames = []
    for decl in body.group(1).split(";"):
        decl = decl.strip()
        if not decl:
            continue
        decl = re.sub(r"^\s*int\s+", "", decl)
        cnames += [re.sub(r"\[.*?\]", "", v).strip() for v in decl.split(",")]
    ftxt = open(os.path.join(ROOT, "fortran", "dlesm_b200_mod.f90")).read()
    fbody = re.search(r"type,\s*bind\(C\)\s*::\s*dlesm_comms(.*?)end type", ftxt, flags=re.S | re.I).group(1)
    fnames = []
    for line in fbody.split("\n"):
        if "::" in line:
            fnames += [v.strip() for v in line.split("::", 1)[1].split("!")[0].split(",") if v.strip()]
    assert fnames == cnames == [n for n, _ in _lib.Comms._fields_], (fnames, cnames)


def test_fortran_calls_pass_as_many_arguments_as_the_interfaces_take():
    """every reference to a dlesm_* interface in fortran/*.f90 and in the lines the patches ADD to the
    reference's field_mod / grid_mod passes exactly the number of arguments its interface declares"""
    import glob
    import re
    arity = {}
    for path in glob.glob(os.path.join(ROOT, "fortran", "*.f90")):
        txt = re.sub(r"&\s*\n\s*&?", " ", open(path).read())
        for fname, args in re.findall(r'(?:function|subroutine)\s+(\w+)\s*\(([^)]*)\)\s*bind\(C', txt, flags=re.I):
            arity[fname.lower()] = len([a for a in args.split(",") if a.strip()])
    assert len(arity) >= 40
    bad, total = [], 0
    files = glob.glob(os.path.join(ROOT, "fortran", "*.f90")) + glob.glob(os.path.join(ROOT, "fortran", "patches", "*.patch"))
    for path in files:
        txt = open(path).read()
        if path.endswith(".patch"):
            txt = "\n".join(l[1:] for l in txt.split("\n") if l.startswith("+") and not l.startswith("+++"))
        txt = "\n".join("" if l.lstrip().startswith("!") else l.split("!")[0] for l in txt.split("\n"))
        txt = re.sub(r"&\s*\n\s*&?", " ", txt)
        for m in re.finditer(r"\b(dlesm_\w+)\s*\(", txt):
            name = m.group(1).lower()
            if name not in arity:
                continue
            i, depth, commas, nonblank = m.end(), 1, 0, False
            while i < len(txt) and depth:
                c = txt[i]
                if c == "(":
                    depth += 1
                elif c == ")":
                    depth -= 1
                elif c == "," and depth == 1:
                    commas += 1
                elif not c.isspace():
                    nonblank = True
                i += 1
            total += 1
            n = commas + 1 if nonblank else 0
            if n != arity[name]:
                bad.append((os.path.basename(path), name, n, arity[name]))
    assert total >= 60
    assert not bad, bad


def test_procedure_pointer_targets_have_the_pointer_interface_characteristics():
    """every `ptr => proc` the patches add to field_mod must name a module procedure of fortran/*.f90 that is NOT
    bind(C) (the reference's abstract interfaces, src/field_mod.f90:66-72,87-93, are not: BIND is a procedure
    characteristic, F2008 12.3.1 / 7.2.2.4) and whose dummy list and declarations equal the abstract interface's"""
    import glob
    import re
    patch = open(os.path.join(ROOT, "fortran", "patches", "field_mod.f90.patch")).read()
    added = "\n".join(l[1:] for l in patch.split("\n") if l.startswith("+") and not l.startswith("+++"))
    targets = dict(re.findall(r"self%(\w+)\s*=>\s*(\w+)", added))
    assert set(targets) == {"read_from_device_c", "write_to_device_c"}
    shim = "\n".join(open(p).read() for p in glob.glob(os.path.join(ROOT, "fortran", "*.f90")))
    shim = re.sub(r"&\s*\n\s*&?", " ", shim)
    # the abstract interface, as the reference has it (both hooks share it)
    want_args = ["from", "to", "startx", "starty", "nx", "ny", "blocking"]
    want_decls = {"type(c_ptr), intent(in), value :: from", "type(c_ptr), intent(in), value :: to",
                  "integer(c_int), intent(in), value :: startx, starty, nx, ny",
                  "logical(c_bool), intent(in), value :: blocking"}
    ref = "/root/reference/finite_difference/src/field_mod.f90"
    if os.path.exists(ref):          # calibrate on the reference itself where it is present
        rt = open(ref).read()
        for comp, iface in (("read_from_device_c", "read_from_device_c_interface"),
                            ("write_to_device_c", "write_to_device_c_interface")):
            m = re.search(r"subroutine\s+%s\s*\(([^)]*)\)([^\n]*)\n(.*?)end subroutine" % iface, rt, flags=re.S | re.I)
            assert m and "bind" not in m.group(2).lower()
            assert [a.strip() for a in m.group(1).split(",")] == want_args
            decl = {" ".join(l.split()) for l in m.group(3).split("\n") if "::" in l}
            assert decl == want_decls, decl
            assert re.search(r"procedure\s*\(\s*%s\s*\)\s*,\s*(?:nopass\s*,\s*pointer|pointer\s*,\s*nopass)\s*::\s*%s\b" % (iface, comp), rt, flags=re.I)
    for comp, proc in targets.items():
        m = re.search(r"^[ \t]*subroutine\s+%s\s*\(([^)]*)\)([^\n]*)\n(.*?)end subroutine" % proc, shim, flags=re.S | re.I | re.M)
        assert m, "%s => %s: no such procedure in fortran/*.f90" % (comp, proc)
        assert "bind" not in m.group(2).lower(), "%s is bind(C); the pointer's interface is not" % proc
        assert [a.strip() for a in m.group(1).split(",")] == want_args
        decl = {" ".join(l.split()) for l in m.group(3).split("\n") if "::" in l}
        assert decl == want_decls, (proc, decl)
        # and it forwards to the C entry point of the same direction
        assert re.search(r"call\s+dlesm_%s\s*\(\s*from\s*,\s*to\s*,\s*startx\s*,\s*starty\s*,\s*nx\s*,\s*ny\s*,\s*blocking\s*\)" % comp,
                         m.group(3), flags=re.I)


def test_fortran_use_statements_name_existing_modules():
    """every `use X` of fortran/*.f90 names a module of the shim or of the reference's src/ (where the
    reference tree is present) or an intrinsic one"""
    import glob
    import re
    ref = "/root/reference/finite_difference/src"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present on this box")
    files = glob.glob(os.path.join(ROOT, "fortran", "*.f90"))
    srcs = files + glob.glob(ref + "/*.f90") + glob.glob(ref + "/*.F90") + glob.glob(ref + "/parallel/*.f90")
    mods = {"iso_c_binding", "iso_fortran_env"}
    for p in srcs:
        mods |= {m.lower() for m in re.findall(r"^\s*module\s+(?!procedure)(\w+)", open(p).read(), flags=re.I | re.M)}
    used = 0
    for p in files:
        for m in re.findall(r"^\s*use\s*(?:,\s*intrinsic\s*::\s*)?(\w+)", open(p).read(), flags=re.I | re.M):
            used += 1
            assert m.lower() in mods, (os.path.basename(p), m)
    assert used >= 5
