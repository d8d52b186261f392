"""The reference's OWN unit tests for the configuration surface ("the
reference's 63 core tests" are the oracle for the boundary) run against this repository's ``zookeeper``
alias package: /root/reference/zookeeper/core/*_test.py are copied (at test time, into a temporary
directory -- never into the repository) and collected by a child pytest whose import path resolves
``zookeeper`` to /root/repo/zookeeper.  Skipped where /root/reference does not exist (the GPU box).
No GPU."""
import glob
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CORE = "/root/reference/zookeeper/core"


@pytest.mark.skipif(not os.path.isdir(REF_CORE), reason="the reference tree is not on this machine")
def test_reference_core_tests_pass_against_the_alias_package(tmp_path):
    files = sorted(glob.glob(os.path.join(REF_CORE, "*_test.py")))
    assert files, "no reference tests found"
    for f in files:
        shutil.copy(f, tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-p", "no:cacheprovider", "."], cwd=tmp_path, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    tail = "\n".join(r.stdout.strip().splitlines()[-15:])
    assert r.returncode == 0, tail
    assert " passed" in tail and "failed" not in tail, tail
