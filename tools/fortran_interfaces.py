"""The C interfaces the reference's Fortran module binds (fortran-bindings/grtcode_fortran.F90:
127-554, `bind(c)` interface blocks), as a fixture, and the checker that compares them with the
prototypes in include/*.h.

No Fortran compiler exists on this machine, so the module cannot be compiled against
libgrtcode.so; what can be checked is that every bound symbol has a prototype here whose
argument list has the same length and, position by position, the same class:

    integer(c_int), value           -> int-like by value (int, Device_t, enum)
    integer(c_int) (by reference)   -> pointer to int-like
    integer(c_int64_t) ...          -> (u)int64 by value / pointer
    real(c_double), value           -> double by value
    real(fp), value                 -> fp_t by value
    real(...) (by reference/array)  -> pointer to double / fp_t
    type(c_ptr), value              -> any object pointer
    type(c_ptr) (array/by ref)      -> pointer to pointer
    character(c_char), dimension(*) -> char pointer

    python tools/fortran_interfaces.py            # regenerate tests/golden/fortran_interfaces.json
                                                  # (needs /root/reference) and check
"""
from __future__ import annotations

import glob
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
F90 = "/root/reference/fortran-bindings/grtcode_fortran.F90"
FIXTURE = os.path.join(ROOT, "tests", "golden", "fortran_interfaces.json")


def parse_f90(path: str) -> dict:
    """{C symbol: {"line": n, "args": [{"name", "class"}], "result": "int"}}."""
    text = open(path).read()
    # join continuation lines
    joined = re.sub(r"&\s*\n\s*", " ", text)
    out = {}
    lineno = {m.group(1).lower(): text[:m.start()].count("\n") + 1
              for m in re.finditer(r"function\s+(\w+)\s*\(", text)}
    for m in re.finditer(r"function\s+(\w+)\s*\(([^)]*)\)\s*result\((\w+)\)\s*bind\(c(?:\s*,\s*name\s*=\s*\"(\w+)\")?\)"
                         r"(.*?)end function", joined, flags=re.S | re.I):
        fname, arglist, result, cname, body = m.groups()
        cname = cname or fname
        names = [a.strip().lower() for a in arglist.split(",") if a.strip()]
        decl = {}
        for line in body.splitlines():
            line = line.split("!")[0].strip()
            if "::" not in line:
                continue
            left, right = line.split("::", 1)
            for var in right.split(","):
                decl[var.strip().lower()] = left.lower()
        args = []
        for n in names:
            d = decl[n]
            value = bool(re.search(r"\bvalue\b", d))
            array = "dimension" in d
            if d.startswith("type(c_ptr)"):
                cls = "ptr" if value and not array else "ptr*"
            elif d.startswith("character"):
                cls = "char*"
            elif d.startswith("integer(kind=c_int64_t)"):
                cls = "int64" if value else "int64*"
            elif d.startswith("integer"):
                cls = "int" if value else "int*"
            elif d.startswith("real(kind=c_double)"):
                cls = "double" if value else "double*"
            elif d.startswith("real(kind=fp)"):
                cls = "fp" if value else "fp*"
            else:
                raise ValueError(f"{cname}: unrecognised declaration {d!r}")
            args.append({"name": n, "class": cls, "optional": "optional" in d})
        rdecl = decl[result.lower()]
        out[cname] = {"line": lineno.get(fname.lower()), "args": args,
                      "result": "int" if rdecl.startswith("integer") else rdecl}
    # subroutines (void functions)
    for m in re.finditer(r"subroutine\s+(\w+)\s*\(([^)]*)\)\s*bind\(c(?:\s*,\s*name\s*=\s*\"(\w+)\")?\)"
                         r"(.*?)end subroutine", joined, flags=re.S | re.I):
        fname, arglist, cname, body = m.groups()
        cname = cname or fname
        names = [a.strip().lower() for a in arglist.split(",") if a.strip()]
        decl = {}
        for line in body.splitlines():
            line = line.split("!")[0].strip()
            if "::" in line:
                left, right = line.split("::", 1)
                for var in right.split(","):
                    decl[var.strip().lower()] = left.lower()
        args = []
        for n in names:
            d = decl[n]
            value = bool(re.search(r"\bvalue\b", d))
            if not d.startswith("integer"):
                raise ValueError(f"{cname}: unrecognised declaration {d!r}")
            args.append({"name": n, "class": "int" if value else "int*", "optional": False})
        out[cname] = {"line": text[:text.lower().find("subroutine " + fname.lower())].count("\n") + 1,
                      "args": args, "result": "void"}
    return out


INTLIKE = {"int", "Device_t", "unsigned", "enum"}


def classify_c(arg: str) -> str:
    a = re.sub(r"\bconst\b", " ", arg)
    a = re.sub(r"\s+", " ", a).strip()
    stars = a.count("*")
    base = a.replace("*", " ").split()
    # drop the parameter name (last token) when a type precedes it
    if len(base) >= 2:
        base = base[:-1]
    t = " ".join(base)
    if stars == 0:
        if t in ("double",):
            return "double"
        if t == "fp_t":
            return "fp"
        if t in ("uint64_t", "int64_t", "size_t"):
            return "int64"
        if t in INTLIKE or t.startswith("enum"):
            return "int"
        return "struct:" + t
    if stars >= 2:
        return "ptr*"
    if t == "char":
        return "char*"
    if t == "double":
        return "double*"
    if t == "fp_t":
        return "fp*"
    if t in ("uint64_t", "int64_t", "size_t"):
        return "int64*"
    if t in INTLIKE:
        return "int*"
    return "ptr"             # pointer to a struct / void


def parse_headers(include_dir: str) -> dict:
    protos = {}
    for h in sorted(glob.glob(os.path.join(include_dir, "*.h"))):
        text = re.sub(r"/\*.*?\*/", " ", open(h).read(), flags=re.S)
        for m in re.finditer(r"EXTERN\s+(?:int|void)\s+(\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
            name, arglist = m.groups()
            args = [a for a in (x.strip() for x in arglist.split(",")) if a and a != "void"]
            protos[name] = {"header": os.path.basename(h), "args": [classify_c(a) for a in args],
                            "text": [re.sub(r"\s+", " ", a) for a in args]}
    return protos


def compatible(f_cls: str, c_cls: str) -> bool:
    if f_cls == c_cls:
        return True
    # a Fortran type(c_ptr) by value carries any object pointer (structs, arrays)
    if f_cls == "ptr" and c_cls.endswith("*"):
        return True
    # an array of c_ptr is a pointer to pointers
    if f_cls == "ptr*" and c_cls in ("ptr*", "ptr"):
        return True
    return False


def check(fixture: dict, protos: dict) -> list[str]:
    problems = []
    for cname, f in sorted(fixture.items()):
        p = protos.get(cname)
        if p is None:
            problems.append(f"{cname}: bound by the Fortran module, no prototype in include/")
            continue
        if len(p["args"]) != len(f["args"]):
            problems.append(f"{cname}: {len(f['args'])} Fortran arguments, {len(p['args'])} in "
                            f"{p['header']}")
            continue
        for k, (fa, ca) in enumerate(zip(f["args"], p["args"])):
            if not compatible(fa["class"], ca):
                problems.append(f"{cname} argument {k + 1} ({fa['name']}): Fortran {fa['class']}, "
                                f"C {ca} ({p['text'][k]})")
    return problems


def main():
    if os.path.exists(F90):
        fx = parse_f90(F90)
        json.dump({"source": "fortran-bindings/grtcode_fortran.F90 (bind(c) interface blocks)",
                   "interfaces": fx}, open(FIXTURE, "w"), indent=1, sort_keys=True)
        print(f"wrote {FIXTURE}: {len(fx)} bound symbols")
    fx = json.load(open(FIXTURE))["interfaces"]
    problems = check(fx, parse_headers(os.path.join(ROOT, "include")))
    for p in problems:
        print("MISMATCH", p)
    print(f"{len(fx)} interfaces checked, {len(problems)} problems")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main())
