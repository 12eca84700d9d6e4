"""The Java binding's JNI shim (java/jni/mp4j_jni.cc) on an image without a JDK: (1) its entry points match the
`native` declarations of com.fenbi.mp4j.comm.Native in name, arity and parameter types (what would otherwise only
show as UnsatisfiedLinkError at run time); (2) the shim runs - master, slaves, host collectives on primitive arrays,
byte transport - against a stand-in JVM made of plain C++ objects (tests/jni_stub/jni.h, tests/cpp/jni_harness.cc)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "java", "jni", "mp4j_jni.cc")
NATIVE = os.path.join(ROOT, "java", "src", "main", "java", "com", "fenbi", "mp4j", "comm", "Native.java")

J2C = {"int": "jint", "long": "jlong", "String": "jstring", "Object": "jobject", "long[]": "jlongArray",
       "byte[]": "jbyteArray", "void": "void"}


def _java_natives():
    out = {}
    for m in re.finditer(r"static native ([\w\[\]]+) (\w+)\(([^)]*)\);", open(NATIVE).read()):
        params = [p.strip().rsplit(" ", 1)[0] for p in m.group(3).split(",") if p.strip()]
        out[m.group(2)] = (J2C[m.group(1)], [J2C[p] for p in params])
    return out


def _shim_functions():
    out = {}
    for m in re.finditer(r"^FN\((\w+), (\w+)\)\(([^)]*)\)", open(SHIM).read(), re.M):
        params = [p.strip().split(" ")[0] for p in m.group(3).split(",")]
        assert params[0] == "JNIEnv*" and params[1] == "jclass", m.group(2)
        out[m.group(2)] = (m.group(1), params[2:])
    return out


def test_native_declarations_match_the_shim():
    java, shim = _java_natives(), _shim_functions()
    assert len(java) >= 30
    assert set(java) == set(shim), (sorted(set(java) - set(shim)), sorted(set(shim) - set(java)))
    for name, sig in java.items():
        assert shim[name] == sig, (name, sig, shim[name])


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_shim_runs_against_a_stand_in_jvm(tmp_path):
    lib = os.path.join(ROOT, "ytk_mp4j_b200", "_lib")
    if not os.path.exists(os.path.join(lib, "libmp4j.so")):
        pytest.skip("libmp4j.so not built")
    exe = str(tmp_path / "jni_harness")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "tests", "jni_stub"),
                    "-I" + os.path.join(ROOT, "include"), SHIM, os.path.join(ROOT, "tests", "cpp", "jni_harness.cc"),
                    "-o", exe, "-L" + lib, "-lmp4j", "-Wl,-rpath," + lib, "-lpthread"], check=True)
    for p in (2, 3):
        r = subprocess.run([exe, str(p)], cwd=str(tmp_path), capture_output=True, text=True, timeout=120)
        assert r.returncode == 0 and "jni harness ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
