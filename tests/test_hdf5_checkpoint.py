"""HDF5 writer/reader and the Keras full-model checkpoint layout."""
import json
import os
import struct

import numpy as np
import pytest

from cori_intml_examples_b200 import nn
from cori_intml_examples_b200.models import mnist, rpv
from cori_intml_examples_b200.utils import hdf5, synthetic
from cori_intml_examples_b200.utils.np_utils import to_categorical


def test_hdf5_round_trip_types_and_nesting(tmp_path):
    p = str(tmp_path / "t.h5")
    with hdf5.File(p, "w") as f:
        f.attrs["title"] = b"hello"
        f.attrs["names"] = [b"a", b"bcd", b"ef"]
        f.attrs["vec"] = np.arange(5, dtype=np.float64)
        g = f.create_group("g1/g2")
        g.create_dataset("w", data=np.arange(24, dtype=np.float32).reshape(2, 3, 4))
        g.create_dataset("it", data=np.asarray(7, dtype=np.int64))
        g.attrs["n"] = np.int32(3)
        many = f.create_group("many")
        for i in range(150):                      # > one symbol-table node (64 entries)
            many.create_dataset("d%03d" % i, data=np.full((2,), i, np.float32))
    with hdf5.File(p, "r") as f:
        assert f.attrs["title"] == b"hello" and [bytes(x) for x in f.attrs["names"]] == [b"a", b"bcd", b"ef"]
        np.testing.assert_array_equal(f.attrs["vec"], np.arange(5.0))
        d = f["g1/g2/w"]
        assert d.shape == (2, 3, 4) and d.dtype == np.float32
        np.testing.assert_array_equal(d[...], np.arange(24, dtype=np.float32).reshape(2, 3, 4))
        np.testing.assert_array_equal(d[:1], np.arange(12, dtype=np.float32).reshape(1, 3, 4))
        assert f["g1/g2/it"][()] == 7 and f["g1"]["g2"].attrs["n"] == 3
        assert len(f["many"].keys()) == 150 and f["many/d149"][...][0] == 149.0
        assert "g1/g2" in f and "nope" not in f
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0                      # superblock version 0
    assert struct.unpack_from("<Q", raw, 40)[0] == len(raw)                      # end-of-file address


def test_keras_checkpoint_layout_and_load_model(tmp_path, fresh_names):
    x, y, _ = synthetic.rpv_like(64, 0)
    m = rpv.build_model((64, 64, 1), [4, 8, 8], [16], dropout=0.2, optimizer="Nadam", lr=0.002, device="cpu")
    m.fit(x, y, batch_size=16, epochs=1, verbose=0)
    p = str(tmp_path / "model_0.h5")
    m.save(p)
    with hdf5.File(p, "r") as f:
        assert f.attrs["keras_version"] == b"2.2.4" and f.attrs["backend"] == b"tensorflow"
        cfg = json.loads(bytes(f.attrs["model_config"]).decode())
        assert cfg["class_name"] == "Model" and cfg["config"]["name"] == "RPVClassifier"
        tc = json.loads(bytes(f.attrs["training_config"]).decode())
        assert tc["loss"] == "binary_crossentropy" and tc["optimizer_config"]["class_name"] == "Nadam"
        names = [bytes(n).decode() for n in f["model_weights"].attrs["layer_names"]]
        assert names[0] == "input_1" and names[1] == "conv2d_1" and names[-1] == "dense_2" and "flatten_1" in names
        g = f["model_weights/conv2d_1"]
        assert [bytes(n) for n in g.attrs["weight_names"]] == [b"conv2d_1/kernel:0", b"conv2d_1/bias:0"]
        assert f["model_weights/conv2d_1/conv2d_1/kernel:0"].shape == (3, 3, 1, 4)                 # HWIO
        assert f["model_weights/dense_1/dense_1/kernel:0"].shape == (8 * 8 * 8, 16)
        ow = [bytes(n).decode() for n in f["optimizer_weights"].attrs["weight_names"]]
        assert ow[0].endswith("iterations:0") and ow[1].endswith("m_schedule:0") and len(ow) == 2 + 2 * 10
        assert f["optimizer_weights"][ow[0]][()] == 4
    m2 = nn.load_model(p, device="cpu")
    np.testing.assert_array_equal(m.predict(x[:8]), m2.predict(x[:8]))
    assert m2.optimizer.class_name == "Nadam" and m2.optimizer.iterations == 4
    assert abs(m2.optimizer.m_schedule - m.optimizer.m_schedule) < 1e-7
    # resumed training continues from the same optimizer state: one more identical step gives identical weights
    m.seed = m2.seed = 5; m.backend.rng.manual_seed(5); m2.backend.rng.manual_seed(5)
    m.train_on_batch(x[:16], y[:16]); m2.train_on_batch(x[:16], y[:16])
    for a, b in zip(m.get_weights(), m2.get_weights()):
        np.testing.assert_allclose(a, b, rtol=0, atol=1e-7)


def test_sequential_checkpoint_and_weights_only(tmp_path):
    xm, ym = synthetic.mnist_like(64, 0)
    m = mnist.build_model(4, 4, 8, 0.1, "Adadelta", device="cpu")
    m.fit(xm, to_categorical(ym, 10), batch_size=32, epochs=1, verbose=0)
    p = str(tmp_path / "seq.h5"); m.save(p)
    m2 = nn.load_model(p, device="cpu")
    assert type(m2).__name__ == "Sequential" and m2.count_params() == m.count_params()
    np.testing.assert_array_equal(m.predict(xm[:4]), m2.predict(xm[:4]))
    wp = str(tmp_path / "w.h5"); m.save_weights(wp)
    m3 = mnist.build_model(4, 4, 8, 0.1, "Adadelta", device="cpu"); m3.load_weights(wp)
    np.testing.assert_array_equal(m.predict(xm[:4]), m3.predict(xm[:4]))


def test_reader_handles_chunked_deflate_dataset(tmp_path):
    """Real RPV files are chunked + gzip-compressed; build such a dataset by hand and read it back."""
    import zlib
    data = np.arange(6 * 4, dtype=np.float32).reshape(6, 4)
    w = hdf5._Writer()
    chunks = []
    for r0 in (0, 4):                                        # chunk shape (4, 4): second chunk is partly outside
        c = np.zeros((4, 4), np.float32); rows = data[r0:r0 + 4]; c[:len(rows)] = rows
        comp = zlib.compress(c.tobytes())
        chunks.append((r0, w.alloc(comp), len(comp)))
    node = bytearray(b"TREE" + struct.pack("<BBHQQ", 1, 0, len(chunks), hdf5.UNDEF, hdf5.UNDEF))
    for r0, addr, size in chunks:
        node += struct.pack("<II", size, 0) + struct.pack("<QQQ", r0, 0, 0) + struct.pack("<Q", addr)
    node += struct.pack("<II", 0, 0) + struct.pack("<QQQ", 8, 0, 0)
    bt = w.alloc(bytes(node))
    layout = struct.pack("<BBB", 3, 2, 3) + struct.pack("<Q", bt) + struct.pack("<III", 4, 4, 4)
    filt = struct.pack("<BBHI", 1, 1, 0, 0) + struct.pack("<HHHH", 1, 0, 1, 1) + struct.pack("<II", 6, 0)
    msgs = [w._msg(0x0001, hdf5._encode_space((6, 4))), w._msg(0x0003, hdf5._encode_dtype(np.float32)),
            w._msg(0x000B, filt), w._msg(0x0008, layout)]
    ds_addr = w.alloc(w._header(msgs))
    root = hdf5._WGroup()

    class _Pre(hdf5._WDataset):
        pass
    # splice the hand-made dataset into the root group
    orig = w.write_dataset
    w.write_dataset = lambda ds: ds_addr if isinstance(ds, _Pre) else orig(ds)
    root.children["comp"] = _Pre(np.zeros(1, np.float32))
    p = str(tmp_path / "c.h5")
    open(p, "wb").write(w.finish(root))
    with hdf5.File(p, "r") as f:
        np.testing.assert_array_equal(f["comp"][...], data)
        np.testing.assert_array_equal(f["comp"][:3], data[:3])


def test_native_cpp_writer_is_byte_identical_to_the_python_writer(tmp_path):
    """csrc/h5writer.cpp (libcibh5.so) is the writer used for checkpoints; the pure-Python writer is its oracle and fall-back."""
    import numpy as np
    from cori_intml_examples_b200.utils import hdf5
    if hdf5.native_lib() is None:
        pytest.skip("libcibh5.so not built")
    root = hdf5._WGroup()
    root.attrs["keras_version"] = "2.2.4"
    root.attrs["backend"] = b"b200"
    root.attrs["model_config"] = '{"class_name": "Sequential", "config": {"name": "x", "layers": []}}' * 20
    g = root.create_group("model_weights")
    g.attrs["layer_names"] = np.array([b"conv2d_1", b"dense_1", b"a_very_long_layer_name_17"])
    rng = np.random.RandomState(0)
    for i in range(70):                                         # > 64 entries: more than one symbol-table node
        d = g.create_dataset("layer_%02d/kernel:0" % i, data=rng.randn(3, 3, 2, 4).astype(np.float32))
        d.attrs["note"] = i
    g.create_dataset("dense_1/dense_1/bias:0", data=np.arange(5, dtype=np.float64))
    root.create_dataset("optimizer_weights/iterations:0", data=np.asarray(7, dtype=np.int64))
    root.create_dataset("empty", data=np.zeros((0, 4), np.float32))
    root.create_dataset("flags", data=np.array([True, False, True]))
    root.create_dataset("names", data=np.array([b"ab", b"cde"]))
    root.create_dataset("u16", data=np.arange(6, dtype=np.uint16).reshape(2, 3))
    py = hdf5._Writer().finish(root)
    cc = hdf5.native_serialize(root)
    assert len(py) == len(cc) and py == cc
    # and the file written through File('w') -- which uses the native writer -- reads back
    path = str(tmp_path / "n.h5")
    with hdf5.File(path, "w") as f:
        f.attrs["a"] = 1.5
        f.create_dataset("g/x", data=np.arange(12, dtype=np.float32).reshape(3, 4))
    with hdf5.File(path, "r") as f:
        assert float(f.attrs["a"]) == 1.5 and np.array_equal(f["g/x"][...], np.arange(12, dtype=np.float32).reshape(3, 4))
    os.environ["CIB_H5_NATIVE"] = "0"
    try:
        hdf5._native[0] = None
        path2 = str(tmp_path / "p.h5")
        with hdf5.File(path2, "w") as f:
            f.attrs["a"] = 1.5
            f.create_dataset("g/x", data=np.arange(12, dtype=np.float32).reshape(3, 4))
        assert open(path, "rb").read() == open(path2, "rb").read()
    finally:
        del os.environ["CIB_H5_NATIVE"]; hdf5._native[0] = None


def test_independent_structural_validator_accepts_checkpoints_and_catches_corruption(tmp_path):
    """utils/h5validate.py is written from the HDF5 format specification and shares no code with the writers: it re-derives every
    structure (superblock, object headers, heaps, B-trees, symbol-table nodes, layouts, attributes) from the bytes.  No h5py /
    libhdf5 exists on the build or GPU machines, so this is the only reader-independent evidence of well-formed files."""
    from cori_intml_examples_b200.utils import h5validate
    m = rpv.build_model((64, 64, 1), [8, 16, 32], [64], dropout=0.5, optimizer="Adam", device="cpu")
    p = str(tmp_path / "ckpt.h5")
    m.save(p)
    rep = h5validate.validate(p)
    assert rep["datasets"] >= 10 + 2 * 10 and rep["groups"] >= 10 and rep["file_bytes"] == os.path.getsize(p)
    blob = bytearray(open(p, "rb").read())
    # a group with more than 2 * leaf-K links (several symbol-table nodes under one B-tree node), an empty group, attributes
    big = str(tmp_path / "big.h5")
    with hdf5.File(big, "w") as f:
        g = f.create_group("g")
        for i in range(150):
            g.create_dataset("d%03d" % i, data=np.arange(i + 1, dtype=np.float32))
        f.create_group("empty")
        f.attrs["note"] = "hello"
    rep = h5validate.validate(big)
    assert rep["datasets"] == 150 and rep["snod_nodes"] >= 3 and rep["groups"] == 3

    def corrupted(mutate):
        b = bytearray(blob)
        mutate(b)
        with pytest.raises(h5validate.H5FormatError):
            h5validate.validate_bytes(bytes(b))

    corrupted(lambda b: b.__setitem__(slice(40, 48), struct.pack("<Q", len(b) + 8)))          # end-of-file address
    corrupted(lambda b: b.__setitem__(13, 4))                                                 # size of offsets
    root = struct.unpack_from("<Q", blob, 64)[0]
    corrupted(lambda b: b.__setitem__(root, 2))                                               # object header version
    corrupted(lambda b: b.__setitem__(slice(root + 2, root + 4), struct.pack("<H", 99)))      # message count
    bt = struct.unpack_from("<Q", blob, 80)[0]
    corrupted(lambda b: b.__setitem__(slice(bt, bt + 4), b"TREX"))                            # B-tree signature
    snod = struct.unpack_from("<Q", blob, bt + 24 + 8)[0]
    assert blob[snod:snod + 4] == b"SNOD" and struct.unpack_from("<H", blob, snod + 6)[0] >= 2

    def swap_first_two_links(b):                                                               # breaks the name order libhdf5 binary-searches
        e0, e1 = bytes(b[snod + 8:snod + 48]), bytes(b[snod + 48:snod + 88])
        b[snod + 8:snod + 48] = e1; b[snod + 48:snod + 88] = e0
    corrupted(swap_first_two_links)


def test_native_and_python_writers_agree_on_empty_groups(tmp_path):
    """Layers without weights (Dropout, Flatten, pooling) are EMPTY groups in a Keras checkpoint: both writers encode them as a B-tree
    root with zero entries and no symbol-table node (what libhdf5 creates); the validator found the earlier non-canonical form."""
    from cori_intml_examples_b200.utils import h5validate
    p = str(tmp_path / "e.h5")
    with hdf5.File(p, "w") as f:
        f.create_group("dropout_1")
        f.create_group("dense_1").create_dataset("kernel:0", data=np.ones((2, 3), np.float32))
    rep = h5validate.validate(p)
    assert rep["groups"] == 3 and rep["snod_nodes"] == 2 and rep["btree_nodes"] == 3
    with hdf5.File(p, "r") as f:
        assert sorted(f.keys()) == ["dense_1", "dropout_1"] and list(f["dropout_1"].keys()) == []


def test_native_reader_matches_the_python_reader(tmp_path, monkeypatch):
    """csrc/h5reader.cpp indexes classic-format files in C++ (the default read path for every checkpoint); the pure-Python reader stays
    for files outside that subset.  Both must present identical trees, attributes, dtypes and values."""
    if hdf5.native_lib() is None or not hasattr(hdf5.native_lib(), "cibh5r_open"):
        pytest.skip("libcibh5.so not built")
    m = mnist.build_model(8, 16, 32, 0.25, "Nadam", device="cpu")
    x, y = synthetic.mnist_like(64, 0)
    m.fit(x, to_categorical(y, 10), batch_size=32, epochs=1, verbose=0)
    p = str(tmp_path / "m.h5")
    m.save(p)
    nat = hdf5.File(p, "r")
    assert nat.reader == "native"
    monkeypatch.setenv("CIB_H5_NATIVE", "0")
    monkeypatch.setattr(hdf5, "_native", [None])
    py = hdf5.File(p, "r")
    assert py.reader == "python"

    def same(a, b, path="/"):
        assert sorted(a.keys()) == sorted(b.keys()), path
        assert set(a.attrs.keys()) == set(b.attrs.keys()), path
        for k in a.attrs:
            assert type(a.attrs[k]) is type(b.attrs[k]) and np.array_equal(a.attrs[k], b.attrs[k]), (path, k)
        n = 0
        for k in a.keys():
            ca, cb = a[k], b[k]
            if hasattr(ca, "keys"):
                n += same(ca, cb, path + k + "/")
            else:
                assert ca.shape == cb.shape and ca.dtype == cb.dtype and np.array_equal(ca[...], cb[...]), path + k
                if ca.shape and ca.shape[0] > 1:
                    assert np.array_equal(ca[1:], cb[1:])
                n += 1
        return n
    assert same(nat._root, py._root) >= 20
    nat.close(); py.close()
    monkeypatch.setattr(hdf5, "_native", [None])
    monkeypatch.delenv("CIB_H5_NATIVE")
    again = nn.load_model(p, device="cpu")                 # the checkpoint path end to end through the native reader
    assert np.allclose(again.predict(x[:8]), m.predict(x[:8]), atol=1e-6)
    bad = tmp_path / "bad.h5"
    bad.write_bytes(b"not an hdf5 file at all" * 10)
    with pytest.raises(Exception):
        hdf5.File(str(bad), "r")
