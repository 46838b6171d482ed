#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — builds the REFERENCE (HyTeG) itself into oracle/_ref/.

Compiles the reference's own source files where they lie under /root/reference with g++ directly
(the reference's CMake build system is NOT run).  The handful of headers / one .cpp its configure
step would generate are hand-written stand-ins under oracle/ref_shim/ (serial, double precision,
generated kernels ON, no MPI / PETSc / ADIOS2 / Trilinos / MPFR).

Outputs (all under oracle/_ref/, git-ignored, shipped to the GPU box with the snapshot):
  obj/*.o                 one object per reference translation unit
  cpu_only/libhyteg_ref.a walberla::core + libhyteg + constant_stencil_operator (P1 part)
  ref_dump                oracle/ref_dump.cpp linked against it: dumps golden fixtures (tests/golden/ref_*.npz
                          are written by oracle/make_ref_fixtures.py from its output) and is the
                          `bench.py --impl reference` CPU arm
  libhyteg_ref_kernels.so the stand-alone generated macro-cell kernels (oracle/Makefile, unchanged)
  meshes/                 the .msh files of the reference's data/meshes/3D (what prependHyTeGMeshDir() resolves to in this build)
  <Test>_ref              an UNMODIFIED reference test (BRIDGE_TESTS), built as it is
  <Test>_b200             the same source with -include integration/hyteg_bridge/swap_P1ConstantLaplaceOperator.hpp: every use of
                          P1ConstantLaplaceOperator becomes the bridge class, linked against hyteg_b200/libhyteg_b200.so
  cpu_only/<Test>_emul    ditto against tests/emul/libhyteg_b200_emul.so (the kernels run sequentially on the host; CPU suite)

Usage: python oracle/build_ref.py [-j N] [--only-driver]
"""
import argparse, os, subprocess, sys, hashlib, concurrent.futures as cf

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("HYTEG_REF", "/root/reference")
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "obj")
CPU = os.path.join(OUT, "cpu_only")  # what the GPU box does not need (listed in .gpurunignore): the archive, the emulation binaries
CXX = os.environ.get("CXX", "g++")
# x86-64-v3 (AVX2+FMA) instead of -march=native: the library travels to another host (the GPU box).
FLAGS = ["-O3", "-march=x86-64-v3", "-std=c++20", "-fPIC", "-w", "-DNDEBUG", "-fopenmp-simd",
         "-I" + os.path.join(HERE, "ref_shim"), "-I" + REF + "/src", "-I" + REF + "/walberla/src",
         "-I" + REF + "/eigen",
         # prependHyTeGMeshDir(): the meshes the reference's own tests name are copied to oracle/_ref/meshes (data, not source), so that
         # an unmodified reference test finds them on the GPU box as well (/root/repo is a link to the snapshot there)
         '-DHYTEG_MESH_DIR="' + os.path.join(OUT, "meshes") + '"']
ROOT = os.path.dirname(HERE)
BRIDGE = os.path.join(ROOT, "integration", "hyteg_bridge")
# unmodified reference tests compiled twice: as they are (the CPU reference) and with P1ConstantLaplaceOperator swapped for the bridge
BRIDGE_TESTS = ["tests/hyteg/convergence/P1GMG3DConvergenceTest.cpp", "tests/hyteg/convergence/P1CG3DConvergenceTest.cpp",
                "tests/hyteg/P1/P1LaplaceOperator3DTest.cpp", "tests/hyteg/convergence/P1GSConvergenceTest.cpp",
                "tests/hyteg/convergence/P1JacobiConvergenceTest.cpp", "tests/hyteg/operators/P1ElementwiseOperatorTest.cpp"]

# sub-trees of the reference that make up the path (walberla::core, libhyteg, the P1 part of
# constant_stencil_operator).  Directories that need packages this image does not have are left out.
SKIP_DIRS = ("/petsc/", "/trilinos/", "/python/", "/checkpointrestore/ADIOS2", "/dataexport/ADIOS2",
             "/adaptiverefinement/", "/experimental/")


ALL = False


def sources():
    src = []
    for root, _, files in os.walk(REF + "/walberla/src/core"):
        src += [os.path.join(root, f) for f in files if f.endswith(".cpp") and not f.endswith(".in.cpp")]
    for root, _, files in os.walk(REF + "/walberla/src/vtk"):
        src += [os.path.join(root, f) for f in files if f.endswith(".cpp")]
    for root, _, files in os.walk(REF + "/src/hyteg"):
        src += [os.path.join(root, f) for f in files if f.endswith(".cpp")]
    cso = REF + "/src/constant_stencil_operator"
    src += [os.path.join(cso, f) for f in ("P1ConstantOperator.cpp",)]
    src += [os.path.join(cso, "P1generatedKernels", f) for f in sorted(os.listdir(cso + "/P1generatedKernels"))
            if f.endswith(".cpp")]
    src.append(os.path.join(HERE, "ref_shim", "waLBerlaBuildInfo.cpp"))
    src = [s for s in sorted(src) if not any(d in s for d in SKIP_DIRS)]
    # oracle/ref_sources.txt: the translation units the driver actually links (read off the linker map, --all ignores it); a fresh
    # build then takes minutes instead of a quarter of an hour
    lst = os.path.join(HERE, "ref_sources.txt")
    if not ALL and os.path.exists(lst):
        need = set(l.strip() for l in open(lst) if l.strip())
        src = [s for s in src if os.path.relpath(s, REF) in need or not s.startswith(REF)]
    return src


def obj_name(s):
    rel = os.path.relpath(s, REF) if s.startswith(REF) else os.path.basename(s)
    return os.path.join(OBJ, hashlib.md5(rel.encode()).hexdigest()[:8] + "_" + os.path.basename(s)[:-4] + ".o")


def compile_one(s):
    o = obj_name(s)
    if os.path.exists(o) and os.path.getmtime(o) >= os.path.getmtime(s):
        return s, 0, ""
    r = subprocess.run([CXX] + FLAGS + ["-c", s, "-o", o], capture_output=True, text=True)
    return s, r.returncode, r.stderr[-2000:]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-j", type=int, default=os.cpu_count())
    ap.add_argument("--only-driver", action="store_true")
    ap.add_argument("--all", action="store_true", help="compile every translation unit of the sub-trees, not only oracle/ref_sources.txt")
    a = ap.parse_args()
    global ALL
    ALL = a.all
    if not os.path.isdir(REF):
        print("reference tree not present (%s): keeping prebuilt oracle/_ref" % REF)
        return 0
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(CPU, exist_ok=True)
    lib = os.path.join(CPU, "libhyteg_ref.a")
    if not a.only_driver or not os.path.exists(lib):
        srcs = sources()
        failed = []
        with cf.ThreadPoolExecutor(a.j) as ex:
            for i, (s, rc, err) in enumerate(ex.map(compile_one, srcs)):
                if rc:
                    failed.append(s)
                    print("FAILED", s, "\n", err, file=sys.stderr)
                elif i % 25 == 0:
                    print("[%d/%d]" % (i, len(srcs)), flush=True)
        if failed:
            print("%d translation units failed (left out of the archive)" % len(failed))
        objs = [obj_name(s) for s in srcs if s not in failed]
        have = set(subprocess.run(["ar", "t", lib], capture_output=True, text=True).stdout.split()) if os.path.exists(lib) else set()
        if have != set(os.path.basename(o) for o in objs) or not fresh(lib, objs):
            if os.path.exists(lib):
                os.remove(lib)
            # one `ar` call per chunk: the command line would be too long otherwise
            for i in range(0, len(objs), 100):
                subprocess.check_call(["ar", "qc", lib] + objs[i:i + 100])
            subprocess.check_call(["ranlib", lib])
    drv = os.path.join(HERE, "ref_dump.cpp")
    if os.path.exists(drv) and not fresh(os.path.join(OUT, "ref_dump"), [drv, os.path.join(HERE, "ref_dump_p2.cpp"), lib]):
        subprocess.check_call([CXX] + FLAGS + [drv, os.path.join(HERE, "ref_dump_p2.cpp"), "-o", os.path.join(OUT, "ref_dump"),
                                               "-Wl,--start-group", lib, "-Wl,--end-group", "-lpthread",
                                               "-Wl,-Map," + os.path.join(OBJ, "ref_dump.map")])
        print("built", os.path.join(OUT, "ref_dump"))
    build_bridge_tests(lib)
    if ALL:
        write_source_list()
    return 0


def write_source_list():
    """oracle/ref_sources.txt <- the archive members the linker actually pulled into ref_dump and the reference tests (their link maps)"""
    import re
    used = set()
    for m in os.listdir(OBJ):
        if m.endswith(".map"):
            used |= set(re.findall(r"libhyteg_ref\.a\(([0-9a-f]{8}_[^)]+\.o)\)", open(os.path.join(OBJ, m)).read()))
    rel = sorted(os.path.relpath(s, REF) for s in sources() if s.startswith(REF) and os.path.basename(obj_name(s)) in used)
    with open(os.path.join(HERE, "ref_sources.txt"), "w") as f:
        f.write("\n".join(rel) + "\n")
    print("oracle/ref_sources.txt: %d translation units" % len(rel))


def fresh(out, deps):
    return os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps if os.path.exists(d))


def build_bridge_tests(lib):
    import shutil
    for sub in ("3D", "2D"):
        os.makedirs(os.path.join(OUT, "meshes", sub), exist_ok=True)
        for f in os.listdir(REF + "/data/meshes/" + sub):
            if f.endswith(".msh") and os.path.getsize(REF + "/data/meshes/" + sub + "/" + f) < (1 << 20):
                shutil.copyfile(REF + "/data/meshes/" + sub + "/" + f, os.path.join(OUT, "meshes", sub, f))
    link = ["-s", "-Wl,--start-group", lib, "-Wl,--end-group", "-lpthread"]  # stripped: the binaries travel to the GPU box with every gpurun call
    for t in BRIDGE_TESTS:
        src, name = os.path.join(REF, t), os.path.basename(t)[:-4]
        inc = ["-I" + REF + "/src/constant_stencil_operator/..", "-I" + REF + "/tests"]
        if not fresh(os.path.join(OUT, name + "_ref"), [src, lib]):
            subprocess.check_call([CXX] + FLAGS + inc + [src, "-o", os.path.join(OUT, name + "_ref")] + link + ["-Wl,-Map," + os.path.join(OBJ, name + ".map")])
        swap = ["-include", os.path.join(BRIDGE, "swap_P1ConstantLaplaceOperator.hpp"), "-I" + BRIDGE, "-I" + os.path.join(ROOT, "include")]
        o = os.path.join(OBJ, name + "_swapped.o")
        hdrs = [os.path.join(BRIDGE, h) for h in os.listdir(BRIDGE)] + [os.path.join(ROOT, "include", "hyteg_b200.h")]
        if not fresh(o, [src] + hdrs):
            subprocess.check_call([CXX] + FLAGS + inc + swap + ["-c", src, "-o", o])
        for tag, d, l in (("_b200", os.path.join(ROOT, "hyteg_b200"), "hyteg_b200"), ("_emul", os.path.join(ROOT, "tests", "emul"), "hyteg_b200_emul")):
            if not os.path.exists(os.path.join(d, "lib" + l + ".so")):
                print("skipping", name + tag, "(lib%s.so not built yet)" % l)
                continue
            exe = os.path.join(CPU if tag == "_emul" else OUT, name + tag)
            rel = os.path.relpath(d, os.path.dirname(exe))
            if fresh(exe, [o, lib, os.path.join(d, "lib" + l + ".so")]):
                continue
            subprocess.check_call([CXX, o, "-o", exe] + link + ["-L" + d, "-l" + l, "-Wl,-rpath,$ORIGIN/" + rel])
            print("built", exe)


if __name__ == "__main__":
    sys.exit(main())
