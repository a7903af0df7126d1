"""The probe image's /probe tree resolves on its own (no docker on this box).

kind-gpu-sim_b200/Dockerfile copies exactly two ELF files into a CUDA *base* image (no toolkit):
/probe/bin/gpu-probe and /probe/lib/libgpuprobe.so.  This test assembles that tree in a scratch
directory, far from the build tree, and checks with an empty LD_LIBRARY_PATH what the image's
loader would have to find: the binary reaches the library through its $ORIGIN rpath, cudart is
linked in statically, and nothing else is needed beyond the C/C++ runtime of the base image
(ubuntu 24.04: glibc 2.39, libstdc++ 3.4.33).  libcuda comes from the NVIDIA container runtime and
libnccl is dlopen'ed (apt-installed by the Dockerfile), so neither may appear as a link-time need."""
import os
import re
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kind-gpu-sim_b200")
BASE_IMAGE_RUNTIME = {"linux-vdso.so.1", "libc.so.6", "libm.so.6", "libstdc++.so.6", "libgcc_s.so.1", "libdl.so.2",
                      "libpthread.so.0", "librt.so.1", "ld-linux-x86-64.so.2"}


def tree(tmp_path):
    probe = tmp_path / "image" / "probe"
    (probe / "bin").mkdir(parents=True)
    (probe / "lib").mkdir()
    shutil.copy2(os.path.join(PKG, "bin", "gpu-probe"), probe / "bin" / "gpu-probe")
    shutil.copy2(os.path.join(PKG, "lib", "libgpuprobe.so"), probe / "lib" / "libgpuprobe.so")
    return probe


def ldd(path, cwd):
    env = {"PATH": "/usr/bin:/bin"}  # no LD_LIBRARY_PATH, as in the image
    out = subprocess.run(["ldd", str(path)], capture_output=True, text=True, env=env, cwd=cwd, check=True).stdout
    deps = {}
    for line in out.splitlines():
        m = re.match(r"\s*(\S+)(?: => (\S+))?", line)
        if m:
            deps[os.path.basename(m.group(1))] = m.group(2) or m.group(1)
    assert "not found" not in out, out
    return deps


def test_dockerfile_copies_exactly_the_two_elf_files_of_the_tree():
    text = open(os.path.join(PKG, "Dockerfile")).read()
    copies = re.findall(r"^COPY (\S+) (\S+)$", text, re.M)
    assert copies == [("kind-gpu-sim_b200/bin/gpu-probe", "/probe/bin/gpu-probe"),
                      ("kind-gpu-sim_b200/lib/libgpuprobe.so", "/probe/lib/libgpuprobe.so")]
    assert "libnccl2" in text and 'ENTRYPOINT ["/probe/bin/gpu-probe"]' in text


def test_relocated_tree_resolves_with_an_empty_library_path(tmp_path):
    probe = tree(tmp_path)
    for elf in (probe / "bin" / "gpu-probe", probe / "lib" / "libgpuprobe.so"):
        deps = ldd(elf, tmp_path)
        for name, where in deps.items():
            if name == "libgpuprobe.so":
                assert os.path.realpath(where) == os.path.realpath(probe / "lib" / "libgpuprobe.so"), where  # via $ORIGIN/../lib
            else:
                assert name in BASE_IMAGE_RUNTIME, f"{elf.name} needs {name}: not in the CUDA base image"
        assert not any(n.startswith(("libcudart", "libcuda", "libnccl", "libprobe_ref")) for n in deps), deps
    dyn = subprocess.run(["readelf", "-d", str(probe / "bin" / "gpu-probe")], capture_output=True, text=True, check=True).stdout
    assert re.search(r"(RUNPATH|RPATH).*\$ORIGIN/\.\./lib", dyn), dyn


def test_symbol_versions_fit_the_base_image(tmp_path):
    probe = tree(tmp_path)
    for elf in (probe / "bin" / "gpu-probe", probe / "lib" / "libgpuprobe.so"):
        syms = subprocess.run(["objdump", "-T", str(elf)], capture_output=True, text=True, check=True).stdout
        glibc = [tuple(int(x) for x in v.split(".")) for v in re.findall(r"GLIBC_(\d+(?:\.\d+)+)", syms)]
        glibcxx = [tuple(int(x) for x in v.split(".")) for v in re.findall(r"GLIBCXX_(\d+(?:\.\d+)+)", syms)]
        assert max(glibc) <= (2, 39), max(glibc)
        assert not glibcxx or max(glibcxx) <= (3, 4, 33), max(glibcxx)


def test_relocated_binary_starts_and_loads_the_library(tmp_path):
    probe = tree(tmp_path)
    r = subprocess.run([str(probe / "bin" / "gpu-probe"), "--bogus"], capture_output=True, text=True,
                       env={"PATH": "/usr/bin:/bin"}, cwd=tmp_path)
    assert r.returncode == 2 and "PROBE FAIL rc=2" in r.stdout  # the dynamic loader found /probe/lib through the rpath
