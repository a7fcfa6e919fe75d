"""Host-side logic of the node-shared host region (no GPU: register=False) and its two-process use."""
import os
import subprocess
import sys

import numpy as np
import pytest

from dynamo_exp_b200.shared_host import SharedHostRegion, split_for_links

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_region_is_page_rounded_and_zeroed():
    with SharedHostRegion.create(5000, register=False) as r:
        assert r.nbytes % 4096 == 0 and r.nbytes >= 5000
        assert r.ptr % 4096 == 0
        a = r.array()
        assert a.shape == (r.nbytes,) and not a.any()
        del a


def test_second_process_sees_and_writes_the_same_bytes():
    r = SharedHostRegion.create(1 << 20, register=False)
    a = r.array()
    a[:] = np.arange(a.size, dtype=np.uint64).astype(np.uint8)
    pid, fd, n = r.handle()
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import numpy as np\n"
        "from dynamo_exp_b200.shared_host import SharedHostRegion\n"
        "r = SharedHostRegion.open((%d, %d, %d), register=False)\n"
        "a = r.array()\n"
        "assert np.array_equal(a, np.arange(a.size, dtype=np.uint64).astype(np.uint8)), 'content'\n"
        "a[100:200] = 7\n"
        "del a; r.close(); print('ok')\n" % (ROOT, pid, fd, n))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stderr
    assert (a[100:200] == 7).all() and a[99] == 99 and a[200] == 200
    del a
    r.close()


def test_open_rejects_a_dead_handle_and_a_short_file():
    with pytest.raises(OSError):
        SharedHostRegion.open((os.getpid(), 10 ** 6, 4096), register=False)
    with SharedHostRegion.create(4096, register=False) as r:
        pid, fd, n = r.handle()
        with pytest.raises(ValueError):
            SharedHostRegion.open((pid, fd, n + 4096), register=False)
    with pytest.raises(ValueError):
        SharedHostRegion.create(0, register=False)


def test_split_for_links():
    assert split_for_links(2) == [(0, 1), (1, 2)]
    assert split_for_links(5) == [(0, 3), (3, 5)]
    assert split_for_links(1) == [(0, 1), (1, 1)]
    assert split_for_links(0) == [(0, 0), (0, 0)]
    assert split_for_links(7, 3) == [(0, 3), (3, 5), (5, 7)]
    with pytest.raises(ValueError):
        split_for_links(3, 0)
