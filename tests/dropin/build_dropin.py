"""Build the DROP-IN proof (test infrastructure, not product):

  tests/dropin/_build/rela.so          the reference's OWN rela/pybind.cc + rela/types.cc,
                                       compiled where they lie under /root/reference, with
                                       rela_b200/cpp/dropin FIRST on the include path so that
                                       "rela/prioritized_replay.h" resolves to the B200 replay.
                                       Actors, thread loops, ModelLocker, Context: unmodified.
  tests/dropin/_build/counting_env.so  a deterministic rela::Env (counting_env.cc), the
                                       stand-in for atari.AtariEnv (no ALE / ROMs here).
  tests/dropin/_build/seq_fixture.so   reference R2D2TransitionBuffer vs rela_b200::SequenceAssembler
                                       (seq_fixture.cc), for tests/test_sequence_assembler.py.
  tests/dropin/_build/test_preplay     the reference's rela/tests/test_prioritized_replay.cc,
                                       compiled as is against the drop-in header.

Only runs where /root/reference exists (the build container); the built files
travel to the GPU box with the snapshot. No reference source is copied.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("RELA_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_build")


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "rela")):
        print("dropin: %s absent, keeping prebuilt %s as is" % (REF, OUT))
        return False
    from rela_b200 import build as b

    b.build_kernels()
    os.makedirs(OUT, exist_ok=True)
    cxx, ld = b._torch_flags()
    cpp = os.path.join(ROOT, "rela_b200", "cpp")
    inc = ["-I" + os.path.join(cpp, "dropin"), "-I" + cpp, "-I" + os.path.join(ROOT, "include"), "-I" + REF]
    # the reference builds with asserts live (rela/CMakeLists.txt:4-6)
    cxx = ["-O3" if f == "-O2" else f for f in cxx if f != "-Wall"] + ["-Wno-register", "-Wno-unused-variable", "-Wno-unused-but-set-variable"]
    hdrs = [os.path.join(dp, f) for dp, _, fs in os.walk(cpp) for f in fs]
    units = {
        "pybind.o": os.path.join(REF, "rela", "pybind.cc"),
        "types.o": os.path.join(REF, "rela", "types.cc"),
        "replay_core.o": os.path.join(cpp, "replay_core.cc"),
        "sharded_core.o": os.path.join(cpp, "sharded_core.cc"),
        "issuer_core.o": os.path.join(cpp, "issuer_core.cc"),
        "counting_env.o": os.path.join(HERE, "counting_env.cc"),
        "seq_fixture.o": os.path.join(HERE, "seq_fixture.cc"),
        "sequence_assembler.o": os.path.join(cpp, "sequence_assembler.cc"),
        "test_preplay.o": os.path.join(REF, "rela", "tests", "test_prioritized_replay.cc"),
    }
    # the UNMODIFIED reference module (its own replay header), for tools/bench_actors.py --impl reference
    ref_obj = os.path.join(OUT, "pybind_ref.o")
    ref_proc = None
    if force or _newer(ref_obj, [units["pybind.o"]]):
        ref_proc = subprocess.Popen(["g++"] + cxx + ["-I" + REF, "-c", units["pybind.o"], "-o", ref_obj])
    procs = []
    for o, src in units.items():
        obj = os.path.join(OUT, o)
        if force or _newer(obj, [src] + hdrs):
            procs.append(subprocess.Popen(["g++"] + cxx + inc + ["-c", src, "-o", obj]))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("drop-in compile failed")
    if ref_proc is not None and ref_proc.wait() != 0:
        raise RuntimeError("reference module compile failed")
    o = lambda n: os.path.join(OUT, n)  # noqa: E731
    os.makedirs(o("reference"), exist_ok=True)
    lib = ["-L" + os.path.join(ROOT, "rela_b200"), "-lrela_b200",
           "-Wl,-rpath," + os.path.join("$ORIGIN", "..", "..", "..", "rela_b200")]
    subprocess.run(["g++", "-shared", "-o", o("rela.so"), o("pybind.o"), o("types.o"), o("replay_core.o"), o("sharded_core.o"), o("issuer_core.o")] + lib + ld,
                   check=True)
    subprocess.run(["g++", "-shared", "-o", o("counting_env.so"), o("counting_env.o")] + ld, check=True)
    subprocess.run(["g++", "-shared", "-o", o("seq_fixture.so"), o("seq_fixture.o"), o("sequence_assembler.o"),
                    o("replay_core.o"), o("sharded_core.o"), o("issuer_core.o"), o("types.o")] + lib + ld, check=True)
    subprocess.run(["g++", "-shared", "-o", os.path.join(o("reference"), "rela.so"), o("pybind_ref.o"), o("types.o")] + ld,
                   check=True)
    import nvidia.cuda_runtime as cr

    ld = ld + ["-L" + os.path.join(list(cr.__path__)[0], "lib"), "-l:libcudart.so.12"]
    pyld = subprocess.run(["python3-config", "--ldflags", "--embed"], capture_output=True, text=True).stdout.split()
    subprocess.run(["g++", "-o", o("test_preplay"), o("test_preplay.o"), o("types.o"), o("replay_core.o"), o("sharded_core.o"), o("issuer_core.o")] + lib + ld
                   + pyld, check=True)
    return True


if __name__ == "__main__":
    build(force="--force" in sys.argv)
    print(sorted(os.listdir(OUT)))
