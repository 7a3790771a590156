"""The JNI boundary text (jni/) against the C ABI, without a JVM.  CPU.

No JDK ships in this image, so the shim cannot be linked or run here; what CAN be checked is that it compiles
against the real include/photon_ml_b200.h (with tests/jni_stub/jni.h standing in for the JDK header), i.e. that
every ABI call it makes has the arguments the header declares, and that the Scala natives and the C functions are
the same set.  ("JNI exports are 1:1 thin wrappers".)
"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
C_FILE = os.path.join(ROOT, "jni", "photon_ml_jni.c")
SCALA = os.path.join(ROOT, "jni", "NativeGLM.scala")


def test_shim_compiles_against_the_abi_header():
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "tests", "jni_stub"),
           "-I", os.path.join(ROOT, "include"), C_FILE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def _c_natives():
    text = open(C_FILE).read()
    return {m.group(2): m.group(1) for m in re.finditer(r"^NATIVE\((\w+),\s*(\w+)\)", text, re.M)}


def _scala_natives():
    text = open(SCALA).read()
    out = {}
    for m in re.finditer(r"@native def (\w+)\((.*?)\)\s*:\s*([\w\[\]]+)", text, re.S):
        out[m.group(1)] = (m.group(2), m.group(3))
    return out


def test_every_scala_native_has_its_c_function_and_back():
    c, s = _c_natives(), _scala_natives()
    assert set(c) == set(s), (sorted(set(c) - set(s)), sorted(set(s) - set(c)))
    jtype = {"Long": "jlong", "Int": "jint", "Double": "jdouble", "Unit": "void", "Array[Byte]": "jbyteArray"}
    for name, (_, ret) in s.items():
        assert c[name] == jtype[ret], (name, ret, c[name])


def test_argument_counts_match():
    text = open(C_FILE).read()
    for name, (args, _) in _scala_natives().items():
        m = re.search(r"NATIVE\(\w+,\s*%s\)\((.*?)\)\s*\{" % name, text, re.S)
        assert m, name
        c_args = [a for a in m.group(1).split(",") if a.strip()]
        scala_args = [a for a in re.split(r",(?![^\[]*\])", args) if a.strip()]
        assert len(c_args) == len(scala_args) + 2, (name, c_args, scala_args)   # + JNIEnv*, jclass


def test_no_elided_bodies_in_the_boundary_text():
    for path in (C_FILE, SCALA):
        text = open(path).read()
        assert "= ..." not in text and "(...)" not in text and "same pattern" not in text, path


def test_shim_only_calls_declared_abi_functions():
    header = open(os.path.join(ROOT, "include", "photon_ml_b200.h")).read()
    declared = set(re.findall(r"\b(pml_\w+)\s*\(", header))
    used = set(re.findall(r"\b(pml_\w+)\s*\(", open(C_FILE).read()))
    assert used <= declared, sorted(used - declared)
    # packing calls must not sit inside a critical region (they allocate and synchronise)
    body = open(C_FILE).read()
    for fn in ("datasetCreate", "batchCreate", "setOffsets", "setWeights", "setNormalization"):
        m = re.search(r"NATIVE\(\w+,\s*%s\)\(.*?\)\s*\{(.*?)\n\}" % fn, body, re.S)
        assert m and "pin(" not in m.group(1) and "GetPrimitiveArrayCritical" not in m.group(1), fn
