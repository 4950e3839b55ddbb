"""include/kfslam_frontend.h is a C header: a strict-C99 translation unit compiles against it, links against the library and runs
the entry points that need no GPU (tests/c/abi_c99.c)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_is_c99_and_library_links(tmp_path):
    import kfslam_b200 as kfs
    kfs.lib()   # builds nothing: fails loudly when the product library is missing
    lib_dir = os.path.join(ROOT, "keyframe-based-slam_b200")
    exe = str(tmp_path / "abi_c99")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "abi_c99.c"), "-L", lib_dir, "-lkfslam_frontend", "-Wl,-rpath," + lib_dir, "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, (out.returncode, out.stdout, out.stderr)
    assert "abi_c99 ok" in out.stdout
