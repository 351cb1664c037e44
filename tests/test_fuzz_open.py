"""CPU: mutation fuzzing of the container validator (tools/fuzz_open.py).  Every mutant of a product-written container is
either rejected by flsgpu_table_open_memory with a status, or accepted - and then the metadata the kernels would follow is
safe to follow: the oracle (which walks the same layout without checks) decodes every column with the image, the decoded
rows and the string heap all ending at PROT_NONE guard pages.  A crash of the worker process fails the test with the seed
image, iteration and mutation (`python tools/fuzz_open.py --replay NAME:ITER` reproduces it)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

import fuzz_open  # noqa: E402


def test_mutated_containers_never_crash_the_open_or_the_walk(tmp_path):
    names = list(fuzz_open.seed_images())
    assert len(names) >= 9  # six numeric types (every encoding each) + the string encodings
    summary, crashes = fuzz_open.run(names, 300, str(tmp_path), chunk=300)
    assert not crashes, crashes[:3]
    for name, s in summary.items():
        assert s["accepted"] + s["rejected"] == 300, name
        assert s["rejected"] > 30 and s["accepted"] > 30, (name, s)  # the mutations reach checked fields and harmless ones
