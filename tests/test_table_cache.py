"""URDF -> flat-table cache on disk (tinyfk_b200/csrc/model.cpp + disk_cache.cpp), no GPU:
a model whose tables come from the cache is the parsed model bit for bit, and a damaged, truncated, foreign or stale file
is never trusted."""
import os
import stat

import numpy as np
import pytest

from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W


@pytest.fixture()
def cache_dir(tmp_path, monkeypatch):
    d = tmp_path / "cache"
    d.mkdir(mode=0o700)
    monkeypatch.setenv("TFK_CACHE_DIR", str(d))
    monkeypatch.delenv("TFK_TABLE_CACHE", raising=False)
    return d


def _same_model(a, b):
    assert (a.n_links, a.n_joints, a.root_link_id) == (b.n_links, b.n_joints, b.root_link_id)
    assert a.get_link_names() == b.get_link_names() and a.get_joint_names() == b.get_joint_names()
    assert np.float64(a.total_mass).tobytes() == np.float64(b.total_mass).tobytes()
    for i in range(a.n_links):
        x, y = a.link_info(i), b.link_info(i)
        assert x[:3] == y[:3] and x[3].tobytes() == y[3].tobytes() and x[4].tobytes() == y[4].tobytes()
    ids = list(range(a.n_joints))
    assert a.get_joint_limits(ids) == b.get_joint_limits(ids)
    assert all(a.is_relevant(l, j) == b.is_relevant(l, j) for l in range(0, a.n_links, 3) for j in range(a.n_joints))


@pytest.mark.parametrize("host_libm", [0, 1])
@pytest.mark.parametrize("urdf", ["pr2.urdf", "kuka.urdf", "fetch.urdf"])
def test_cached_tables_equal_the_parsed_model(cache_dir, urdf, host_libm):
    x = W.urdf_text(urdf)
    parsed = capi.Model(x, host_libm=host_libm)
    assert not parsed.tables_from_cache()
    files = sorted(os.listdir(cache_dir))
    assert len(files) == 1 and files[0].startswith("tables_") and files[0].endswith(".tfkt")
    cached = capi.Model(x, host_libm=host_libm)
    assert cached.tables_from_cache()
    _same_model(parsed, cached)
    # the other host libm is another key (the origin quaternions depend on it)
    other = capi.Model(x, host_libm=1 - host_libm)
    assert not other.tables_from_cache() and len(os.listdir(cache_dir)) == 2
    # structure edits and state behave the same on a model that never saw the XML
    pid = cached.get_link_ids([cached.get_link_names()[3]])[0]
    assert cached.add_new_link("added", pid, [0.1, 0.2, 0.3], [0.3, 0.2, 0.1]) == parsed.add_new_link("added", pid, [0.1, 0.2, 0.3], [0.3, 0.2, 0.1])
    _same_model(parsed, cached)
    with pytest.raises(ValueError):
        cached.add_new_link("added", pid, [0, 0, 0], [0, 0, 0])
    j = list(range(cached.n_joints))
    cached.set_joint_angles(j, np.linspace(-1, 1, len(j)))
    assert np.array_equal(cached.get_joint_angles(j), np.linspace(-1, 1, len(j)))


def test_com_links_and_mass_survive_the_cache(cache_dir):
    """the "<name>_com" links (src/tinyfk.cpp:89-111) and their inertials are part of the tables: the centre-of-mass plan of a
    cached model lists the same links"""
    x = W.urdf_text("fetch.urdf")
    a, b = capi.Model(x), capi.Model(x)
    assert not a.tables_from_cache() and b.tables_from_cache()
    na, nb = a.get_link_names(), b.get_link_names()
    assert [n for n in na if n.endswith("_com")] == [n for n in nb if n.endswith("_com")] and len([n for n in nb if n.endswith("_com")]) == 20
    assert a.total_mass == b.total_mass > 0


def _only_file(d):
    (f,) = os.listdir(d)
    return os.path.join(d, f)


@pytest.mark.parametrize("damage", ["flip_payload", "flip_key", "truncate", "magic", "append", "empty"])
def test_damaged_files_are_rebuilt_not_trusted(cache_dir, damage):
    x = W.urdf_text("kuka.urdf")
    ref = capi.Model(x)
    path = _only_file(cache_dir)
    blob = bytearray(open(path, "rb").read())
    if damage == "flip_payload":
        blob[len(blob) // 2] ^= 0x40
    elif damage == "flip_key":
        blob[9] ^= 1
    elif damage == "truncate":
        blob = blob[: len(blob) - 17]
    elif damage == "magic":
        blob[0:8] = b"TFKCUBN1"  # the cubin cache's magic: a foreign file under the right name
    elif damage == "append":
        blob += b"\0" * 8
    else:
        blob = bytearray()
    open(path, "wb").write(bytes(blob))
    m = capi.Model(x)
    assert not m.tables_from_cache()
    _same_model(ref, m)
    assert capi.Model(x).tables_from_cache()  # ... and the parse wrote a good file again


def test_consistent_header_with_bad_tables_is_rejected(cache_dir):
    """a file whose header verifies but whose payload holds out-of-range indices (written by a buggy or hostile producer that
    knows the format) must fail the bounds checks of model.cpp:load_tables, not index out of range"""
    import struct

    x = W.urdf_text("kuka.urdf")
    ref = capi.Model(x)
    path = _only_file(cache_dir)
    blob = bytearray(open(path, "rb").read())
    hdr = 8 + 4 * 8
    payload = blob[hdr:]

    def fnv(b):
        h = 1469598103934665603
        for c in b:
            h = ((h ^ c) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        return h

    for off, val in ((12, 1 << 19), (4, 1 << 30), (12, -5)):  # root far out of range, absurd link count, negative root
        p = bytearray(payload)
        p[off:off + 4] = struct.pack("<i", val)
        blob[hdr:] = p
        blob[8 + 24:8 + 32] = struct.pack("<Q", fnv(p))
        open(path, "wb").write(bytes(blob))
        m = capi.Model(x)
        assert not m.tables_from_cache()
        _same_model(ref, m)


def test_switch_and_untrusted_directory(cache_dir, monkeypatch):
    x = W.urdf_text("kuka.urdf")
    monkeypatch.setenv("TFK_TABLE_CACHE", "0")
    capi.Model(x)
    assert os.listdir(cache_dir) == [] and not capi.Model(x).tables_from_cache()
    monkeypatch.delenv("TFK_TABLE_CACHE")
    os.chmod(cache_dir, 0o777)  # writable by others: neither read nor written
    capi.Model(x)
    assert os.listdir(cache_dir) == [] and not capi.Model(x).tables_from_cache()
    os.chmod(cache_dir, stat.S_IRWXU)
    capi.Model(x)
    assert capi.Model(x).tables_from_cache()
    # malformed input never leaves a file behind
    n = len(os.listdir(cache_dir))
    with pytest.raises(ValueError):
        capi.Model("<robot name='r'><link name='a'/><link name='a'/></robot>")
    assert len(os.listdir(cache_dir)) == n


def test_concurrent_processes_share_one_cache(cache_dir):
    """several processes that meet the same URDF at the same time (one rank per GPU under torchrun): every one gets a correct
    model, whichever of them writes the file (temporary + rename), and the file they leave behind verifies"""
    import subprocess
    import sys

    code = ("import sys; sys.path.insert(0, %r)\n"
            "from tinyfk_b200 import capi, workloads as W\n"
            "m = capi.Model(W.urdf_text('pr2.urdf'))\n"
            "print(m.n_links, m.n_joints, int(m.tables_from_cache()))\n" % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    procs = [subprocess.Popen([sys.executable, "-c", code], stdout=subprocess.PIPE, text=True, env=dict(os.environ)) for _ in range(6)]
    outs = [p.communicate()[0].split() for p in procs]
    assert all(p.returncode == 0 for p in procs)
    assert all(o[:2] == ["162", "45"] for o in outs)
    files = os.listdir(cache_dir)
    assert len(files) == 1 and not files[0].endswith(".tmp")
    assert capi.Model(W.urdf_text("pr2.urdf")).tables_from_cache()
