"""bench.py contract checks that need no GPU: the reference arm (CPU port of the reference's path on
the host cores) prints one JSON line with the agreed keys, and the work model used for the roofline
matches SURVEY.md 8(d)."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def test_work_model_matches_survey():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.modmuls_per_perm(3, 8, 57, 5) == 828          # C2
    assert bench.modmuls_per_perm(5, 8, 60, 5) == 2000         # C3
    assert bench.modmuls_per_perm(3, 8, 29, 17) == 598         # C1 / C5
    assert bench.modmuls_per_perm(3, 8, 31, 17) == 626         # KAT-pinned default set
    assert 828 * bench.IMUL_PER_MODMUL == 112608


def test_reference_arm_json_line():
    env = dict(os.environ, OMP_NUM_THREADS="1")   # torchrun exports this; the arm must not obey it
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--cpu-seconds", "1"], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "poseidon_permutations_per_second"
    assert line["unit"] == "perm/s" and line["higher_is_better"] is True and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "perm/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in line["config"] and line["config"]["workload"].startswith("c2")


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_executed_work_block_reads_the_build_time_kernel_facts():
    """bench.py's roofline.exec block: the kernel a context reports is looked up in the SASS facts written at build
    time (sponge_b200/lib/kernel_sass.json); no GPU needed for the lookup itself."""
    sys.path.insert(0, ROOT)
    import bench

    class FakeCtx:
        def __init__(self, name):
            self.name = name

        def describe(self):
            return dict(kernel_name=self.name, constants_home="bank", variant=5, image_bytes=0, image_shared_by=1)

    for name in ("sponge_kernel<bn254,t=3,alpha=5,variant=5>", "sponge_kernel<bn254,t=3,alpha=5,variant=6,smem>",
                 "sponge_kernel<bn254,t=5,alpha=5,variant=5,smem>", "sponge_kernel<bls12_381,t=3,alpha=17,variant=3>",
                 "sponge_kernel<bn254,t=3,alpha=5,variant=5,ragged>"):
        info, traffic = bench.executed_work(FakeCtx(name), "c2", 24, 1 << 24, 96.0, dict(sm_mhz=1965.0))
        assert info["kernel"] == name
        assert "sass_sha256" in info and info["static_instructions"] > 1000, (name, info)
