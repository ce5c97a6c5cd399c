"""The JNI glue ships as source (this image has no JDK).  What can be checked here: the C file compiles
against the C ABI header with a stand-in jni.h (types and signatures agree with include/oryx_b200.h), and
every native method of OryxB200.java has exactly one JNI function and vice versa."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_jni_glue_compiles_against_the_abi_header():
    subprocess.check_call(["gcc", "-fsyntax-only", "-Wall", "-Wextra", "-Werror",
                           "-I" + os.path.join(ROOT, "tests", "cpp", "jni_stub"), "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "jni", "oryx_b200_jni.c")])


def test_every_native_method_has_its_jni_function():
    java = open(os.path.join(ROOT, "jni", "OryxB200.java")).read()
    c = open(os.path.join(ROOT, "jni", "oryx_b200_jni.c")).read()
    natives = set(re.findall(r"public static native [\w\[\]\.]+\s+(\w+)\(", java))
    funcs = set(re.findall(r"JNICALL FN\((\w+)\)", c))
    assert natives and natives == funcs, (sorted(natives - funcs), sorted(funcs - natives))
