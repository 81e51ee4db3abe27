#!/usr/bin/env python
"""Generates fortran/fempar_b200_bindings.f90 -- the ISO_C_BINDING interface module of EVERY entry point declared in
include/fempar_b200.h -- from the header itself, so that the two cannot drift apart (tests/test_fortran_bindings.py checks it).

Mapping of the C types of the header:
  opaque handles (fb2_space_t, fb2_matrix_t, fb2_vector_t)  -> type(c_ptr), value          (handle*  -> type(c_ptr), by reference)
  int / int32_t / uint32_t / int64_t / double by value      -> integer(c_int|c_int32_t|c_int64_t) / real(c_double), value
  pointers to scalars or arrays of those                    -> assumed-size arrays of the same kind (a scalar actual argument
                                                               is passed by reference, sequence association)
  const char*                                               -> character(kind=c_char) :: name(*)
  void* / handle arrays                                     -> type(c_ptr), value
  struct pointers                                           -> type(<struct>), by reference
usage: python fortran/gen_bindings.py [--check]
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "fempar_b200.h")
OUT = os.path.join(ROOT, "fortran", "fempar_b200_bindings.f90")

HANDLES = {"fb2_space_t", "fb2_matrix_t", "fb2_vector_t"}
SCALARS = {"int": "integer(c_int)", "int32_t": "integer(c_int32_t)", "uint32_t": "integer(c_int32_t)",
           "int64_t": "integer(c_int64_t)", "double": "real(c_double)", "unsigned": "integer(c_int)"}
STRUCTS = {"fb2_ils_params_t", "fb2_ils_ext_params_t"}
# pointer arguments the header documents as "may be NULL": type(c_ptr), value (pass c_loc(array) or c_null_ptr)
NULLABLE = {"history", "history_extra", "func_params", "mat_params", "fixed_dof_values", "elmat", "elvec", "val", "weight",
            "host_in", "ms_total", "ms_spmv", "irp_host", "ja_host", "gid_out", "part_out", "dinv"}
# DEVICE pointers (what fb2_vec_device_ptr returns): a Fortran array dummy would pass a host address, so type(c_ptr), value too
DEVICE_PTRS = {"x_dev", "y_dev"}


def strip_comments(txt):
    txt = re.sub(r"/\*.*?\*/", " ", txt, flags=re.S)
    return re.sub(r"//[^\n]*", " ", txt)


def prototypes(txt):
    body = strip_comments(txt)
    body = re.sub(r"#[^\n]*", " ", body)
    body = re.sub(r"typedef\s+struct\s*\{.*?\}\s*\w+\s*;", " ", body, flags=re.S)
    body = re.sub(r"typedef[^;]*;", " ", body)
    body = body.replace('extern "C" {', " ").replace("}", " ")
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(fb2_\w+)\s*\(([^;{]*)\)\s*;", body, flags=re.S):
        ret, name, args = " ".join(m.group(1).split()), m.group(2), " ".join(m.group(3).split())
        yield ret, name, args


def parse_arg(a):
    a = a.strip()
    if a in ("void", ""):
        return None
    arr = re.search(r"\[(\d*)\]$", a)
    if arr:
        a = a[:arr.start()].strip()
    const = bool(re.search(r"\bconst\b", a))
    a = re.sub(r"\bconst\b", " ", a)
    m = re.match(r"(.*?)(\w+)$", a.strip())
    ctype, name = m.group(1).strip(), m.group(2)
    nptr = ctype.count("*") + (1 if arr else 0)
    base = ctype.replace("*", "").strip()
    return dict(base=base, nptr=nptr, name=name, const=const, dim=(arr.group(1) if arr else None))


def fortran_decl(a, fname=""):
    b, n, name = a["base"], a["nptr"], a["name"]
    intent = "intent(in)" if a["const"] else "intent(inout)"
    if b in HANDLES:
        if n == 0:
            return "type(c_ptr), value :: %s" % name
        # handle*: the new handle of a create call (scalar by reference), or the handle list of a *_multi call (array)
        return "type(c_ptr) :: %s%s" % (name, "(*)" if fname.endswith("_multi") else "")
    if (b == "void" and n >= 1) or (n >= 1 and (name in NULLABLE or name in DEVICE_PTRS) and b in SCALARS):
        return "type(c_ptr), value :: %s" % name
    if b == "char" and n == 1:
        return "character(kind=c_char), %s :: %s(*)" % (intent, name)
    if b in STRUCTS:
        return "type(%s), %s :: %s" % (b, intent, name)
    if b in SCALARS:
        if n == 0:
            return "%s, value :: %s" % (SCALARS[b], name)
        shape = "(%s)" % a["dim"] if a["dim"] else "(*)"
        return "%s, %s :: %s%s" % (SCALARS[b], intent, name, shape)
    raise ValueError("unmapped C type %r in argument %r" % (b, name))


def struct_defs(txt):
    out = []
    for m in re.finditer(r"typedef\s+struct\s*\{(.*?)\}\s*(\w+)\s*;", strip_comments(txt), flags=re.S):
        fields, name = m.group(1), m.group(2)
        out.append("  type, bind(C) :: %s" % name)
        for f in fields.split(";"):
            f = " ".join(f.split())
            if not f:
                continue
            ctype, names = f.split(" ", 1)
            out.append("     %s :: %s" % (SCALARS[ctype], names))
        out.append("  end type %s" % name)
    return out


def constants(txt):
    out = []
    for m in re.finditer(r"#define\s+(FB2_\w+)\s+(-?\d+)\b", txt):
        out.append("  integer(c_int), parameter :: %s = %s" % (m.group(1), m.group(2)))
    return out


def generate():
    txt = open(HDR).read()
    L = ["! GENERATED by fortran/gen_bindings.py from include/fempar_b200.h -- do not edit by hand.",
         "! ISO_C_BINDING interfaces of every entry point of libfempar_b200.so.  Optional C arguments (documented as \"may be NULL\"",
         "! in the header) are passed as c_null_ptr through the type(c_ptr), value dummies, or with c_loc of the actual array.",
         "module fempar_b200_bindings",
         "  use iso_c_binding",
         "  implicit none",
         "  public"]
    L += constants(txt)
    L += struct_defs(txt)
    L.append("  interface")
    names = []
    for ret, name, args in prototypes(txt):
        names.append(name)
        parsed = [parse_arg(a) for a in split_args(args)]
        parsed = [p for p in parsed if p]
        argn = ", ".join(p["name"] for p in parsed)
        if ret.replace(" ", "") in ("constchar*", "void*"):
            head = "     function %s(%s) bind(C, name='%s') result(res)" % (name, argn, name)
            tail = ["       type(c_ptr) :: res"]
            end = "     end function %s" % name
        else:
            head = "     integer(c_int) function %s(%s) bind(C, name='%s')" % (name, argn, name)
            tail = []
            end = "     end function %s" % name
        L += wrap(head)
        L.append("       import")
        for p in parsed:
            L.append("       " + fortran_decl(p, name))
        L += tail
        L.append(end)
    L.append("  end interface")
    L.append("end module fempar_b200_bindings")
    return "\n".join(L) + "\n", names


def split_args(args):
    return [a for a in args.split(",")]


def wrap(line, width=128):
    out = []
    while len(line) > width:
        cut = line.rfind(",", 0, width - 2)
        out.append(line[:cut + 1] + " &")
        line = "          " + line[cut + 1:].lstrip()
    out.append(line)
    return out


if __name__ == "__main__":
    text, names = generate()
    if "--check" in sys.argv:
        ok = os.path.exists(OUT) and open(OUT).read() == text
        print("fortran/fempar_b200_bindings.f90 is %s (%d entry points)" % ("up to date" if ok else "STALE", len(names)))
        sys.exit(0 if ok else 1)
    open(OUT, "w").write(text)
    print("wrote %s: %d entry points" % (OUT, len(names)))
