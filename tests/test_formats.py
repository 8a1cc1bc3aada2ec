"""On-disk formats either side of the stepping path: PNG heightmaps and model-tree YAML.

CPU-only: the readers are host code of the C-ABI library (no kernel is launched).  Golden data comes from the reference
itself (tools/make_golden.py): heightmap levels decoded from its env_data/*.png, trees fitted/saved/evaluated by its own
scripts/ModelTree/model_tree.py.
"""
import os
import struct
import zlib

import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import TfodeError, abi, load_heightmap_png, make_config, parse_model_tree, reference_heightmap

yaml = pytest.importorskip("yaml")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---------------------------------------------------------------------------------------------------------------------- PNG
def _chunk(t, body):
    return struct.pack(">I", len(body)) + t + body + struct.pack(">I", zlib.crc32(t + body) & 0xFFFFFFFF)


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def encode_png(rows, width, depth, ctype, filters, palette=None, idat_split=1, interlace=0, height=None):
    """rows: list of packed scanlines (bytes).  filters: filter type per row (cycled)."""
    channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    bpp = max(1, channels * depth // 8)
    raw = bytearray()
    prev = bytes(len(rows[0]))
    for y, line in enumerate(rows):
        f = filters[y % len(filters)]
        out = bytearray(len(line))
        for i, v in enumerate(line):
            a = line[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            pred = (0, a, b, (a + b) >> 1, _paeth(a, b, c))[f]
            out[i] = (v - pred) & 255
        raw += bytes([f]) + out
        prev = line
    z = zlib.compress(bytes(raw), 6)
    png = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", width, height or len(rows), depth, ctype, 0, 0, interlace))
    png += _chunk(b"pHYs", struct.pack(">IIB", 2835, 2835, 1))  # an ancillary chunk, like the reference's files carry
    if palette is not None:
        png += _chunk(b"PLTE", bytes(palette))
    step = max(1, (len(z) + idat_split - 1) // idat_split)
    for k in range(0, len(z), step):
        png += _chunk(b"IDAT", z[k:k + step])
    return png + _chunk(b"IEND", b"")


def _levels(h=37, w=53, seed=3):
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    return ((np.sin(xx / 5.0) * 60 + np.cos(yy / 7.0) * 50 + 128 + rng.randint(-8, 9, size=(h, w))).clip(0, 255)).astype(np.uint8)


@pytest.mark.parametrize("ctype", [0, 2, 4, 6])
@pytest.mark.parametrize("filters", [[0], [1], [2], [3], [4], [0, 1, 2, 3, 4]])
def test_png_8bit_colour_types_and_filters(ctype, filters):
    lv = _levels()
    channels = {0: 1, 2: 3, 4: 2, 6: 4}[ctype]
    rng = np.random.RandomState(5)
    px = rng.randint(0, 256, size=lv.shape + (channels,)).astype(np.uint8)
    px[..., 0] = lv   # the reference keeps the first byte of each pixel (ode/heightfield.hh:54)
    rows = [px[y].tobytes() for y in range(lv.shape[0])]
    h = load_heightmap_png(encode_png(rows, lv.shape[1], 8, ctype, filters, idat_split=3))
    assert h.shape == lv.shape and h.dtype == np.float32
    assert np.array_equal(h, (lv.astype(np.float64) * 0.3 / 255).astype(np.float32))


@pytest.mark.parametrize("depth", [1, 2, 4])
def test_png_low_bit_depth_grey_and_palette(depth):
    hgt, wid = 11, 29   # width not a multiple of 8/depth: the last byte of a scanline is partly padding
    rng = np.random.RandomState(depth)
    idx = rng.randint(0, 1 << depth, size=(hgt, wid))
    rows = []
    for y in range(hgt):
        bits = "".join(format(v, "0%db" % depth) for v in idx[y])
        bits += "0" * (-len(bits) % 8)
        rows.append(bytes(int(bits[i:i + 8], 2) for i in range(0, len(bits), 8)))
    grey = load_heightmap_png(encode_png(rows, wid, depth, 0, [0, 2, 4]))
    want = (idx * 255 // ((1 << depth) - 1)).astype(np.float64) * 0.3 / 255
    assert np.array_equal(grey, want.astype(np.float32))
    pal = rng.randint(0, 256, size=((1 << depth), 3)).astype(np.uint8)
    got = load_heightmap_png(encode_png(rows, wid, depth, 3, [1, 3], palette=pal.ravel().tolist()))
    assert np.array_equal(got, (pal[idx, 0].astype(np.float64) * 0.3 / 255).astype(np.float32))


def test_png_z_scale_and_file_path(tmp_path):
    lv = _levels(8, 8)
    png = encode_png([lv[y].tobytes() for y in range(8)], 8, 8, 0, [4])
    p = tmp_path / "m.png"
    p.write_bytes(png)
    assert np.array_equal(load_heightmap_png(str(p), 1.0), (lv.astype(np.float64) * 1.0 / 255).astype(np.float32))


def test_png_errors_are_loud(tmp_path):
    lv = _levels(8, 8)
    rows = [lv[y].tobytes() for y in range(8)]
    good = encode_png(rows, 8, 8, 0, [1])
    with pytest.raises(TfodeError, match="Can't open"):   # the reference's message, ode/heightfield.hh:48
        load_heightmap_png(str(tmp_path / "missing.png"))
    with pytest.raises(TfodeError, match="signature"):
        load_heightmap_png(b"JFIF" + good[4:])
    bad = bytearray(good); bad[40] ^= 0x55
    with pytest.raises(TfodeError, match="CRC|inflate|past the end"):
        load_heightmap_png(bytes(bad))
    with pytest.raises(TfodeError, match="truncated|past the end"):
        load_heightmap_png(good[:-20])
    with pytest.raises(TfodeError, match="interlaced"):
        load_heightmap_png(encode_png(rows, 8, 8, 0, [0], interlace=1))
    with pytest.raises(TfodeError, match="16-bit"):
        load_heightmap_png(encode_png([r + r for r in rows], 8, 16, 0, [0]))
    with pytest.raises(TfodeError, match="inflate"):
        load_heightmap_png(encode_png(rows[:5], 8, 8, 0, [0], height=8))


def test_reference_heightmaps_match_the_committed_levels():
    """The decoder on the reference's own files (only where /root/reference exists) against the PIL-decoded fixture."""
    root = "/root/reference/env_data"
    if not os.path.isdir(root):
        pytest.skip("/root/reference is not on this machine")
    for kind, shape in (("rock_step", (200, 200)), ("rock_groove", (200, 200)), ("rock_groove_large", (400, 400))):
        h = load_heightmap_png(os.path.join(root, "heightmap_%s.png" % kind))
        assert h.shape == shape
        assert np.array_equal(h, reference_heightmap(kind))


def test_committed_heightmap_levels_have_the_reference_statistics():
    lv = dict(np.load(os.path.join(os.path.dirname(GOLD), os.pardir, "tf_ode_b200", "data", "heightmap_levels.npz")))
    assert lv["rock_step"].shape == (200, 200) and lv["rock_step"].min() == 0 and lv["rock_step"].max() == 254
    assert lv["rock_groove"].shape == (200, 200) and lv["rock_groove"].min() == 86 and lv["rock_groove"].max() == 255
    assert lv["rock_groove_large"].shape == (400, 400) and (lv["rock_groove_large"][0] == 255).all()
    assert abs(float(reference_heightmap("rock_groove").max()) - 0.3) < 1e-7


# ---------------------------------------------------------------------------------------------------------------------- YAML trees
def _flatten_like_the_loader(d):
    """Linear_model_tree::_build_tree_recursively (model_tree.cc:21-45) on the PyYAML dict."""
    nodes, params = [], []
    count = [1]

    def rec(node_id):
        e = d[node_id]
        p = e["model params"] if e["terminal"] else e["split params"]
        k = len(nodes)
        nodes.append(dict(id=node_id, terminal=bool(e["terminal"]), child=[-1, -1], param_off=len(params), n_params=len(p)))
        params.extend(float(x) for x in p)
        if not e["terminal"]:
            for c in (0, 1):
                count[0] += 1
                nodes[k]["child"][c] = len(nodes)
                rec(count[0])
    rec(1)
    return nodes, np.array(params)


@pytest.mark.parametrize("name", ["1", "2", "poly2"])
def test_yaml_parser_matches_pyyaml_on_the_reference_written_trees(name):
    text = open(os.path.join(GOLD, "model_tree_%s.yaml" % name)).read()
    nodes, par = parse_model_tree(text)
    wn, wp = _flatten_like_the_loader(yaml.safe_load(text))
    assert len(nodes) == len(wn)
    for a, b in zip(nodes, wn):
        assert a["id"] == b["id"] and a["terminal"] == b["terminal"] and list(a["child"]) == b["child"]
        assert a["param_off"] == b["param_off"] and a["n_params"] == b["n_params"]
    assert np.array_equal(par, wp)   # strtod and PyYAML's float() agree bit for bit


def test_yaml_parser_accepts_flow_style_quotes_and_comments():
    d = yaml.safe_load(open(os.path.join(GOLD, "model_tree_1.yaml")).read())
    for style in (True, None):   # all-flow leaves (PyYAML < 5.1 default) and mixed
        text = "# header comment\n" + yaml.dump(d, default_flow_style=style, sort_keys=False, width=60) if style is None else None
        if text is None:
            lines = ["# Straight Model Tree"]
            for k, v in d.items():
                key = "model params" if v["terminal"] else "split params"
                lines += ["%d:   # node" % k, "  'info': \"%s\"" % v["info"], "  terminal: %s" % ("True" if v["terminal"] else "no"),
                          "  \"%s\": [%s]" % (key, ",\n     ".join(repr(float(x)) for x in v[key]))]
            text = "\n".join(lines) + "\n"
        nodes, par = parse_model_tree(text)
        wn, wp = _flatten_like_the_loader(d)
        assert [n["id"] for n in nodes] == [n["id"] for n in wn]
        assert np.array_equal(par, wp)


def test_yaml_parser_errors_follow_the_reference():
    with pytest.raises(TfodeError, match="Read 3 split parameters for the node 1 instead of the 2 required for a straight tree"):
        parse_model_tree("1:\n  terminal: false\n  split params: [0, 1.0, 2.0]\n")   # model_tree.cc:33-35
    nodes, par = parse_model_tree("1:\n  terminal: false\n  split params: [0.5, 1.0, 2.0]\n2:\n  terminal: true\n  model params: [1, 2, 3]\n"
                                  "3:\n  terminal: true\n  model params: [4, 5, 6]\n", oblique=True)
    assert [n["id"] for n in nodes] == [1, 2, 3] and par.tolist() == [0.5, 1, 2, 1, 2, 3, 4, 5, 6]
    with pytest.raises(TfodeError, match="node 3"):
        parse_model_tree("1:\n  terminal: false\n  split params: [0, 1.0]\n2:\n  terminal: true\n  model params: [1, 2]\n")
    with pytest.raises(TfodeError, match="not a number"):
        parse_model_tree("1:\n  terminal: true\n  model params: [1, abc]\n")
    with pytest.raises(TfodeError, match="no 'terminal'"):
        parse_model_tree("1:\n  model params: [1, 2]\n")


# ---------------------------------------------------------------------------------------------------------------------- oracle trees vs the reference's Python
@pytest.fixture(scope="module")
def tree_oracle():
    o = oracle_lib.Oracle(make_config(abi.PRESET_SCENE_1, 1, scene=abi.SCENE_PLANE), "double")
    yield o
    o.close()


def test_oracle_model_tree_matches_reference_python_predictions(tree_oracle):
    """model_tree.py's predict() (sklearn dot product) vs the restated C++ traversal: same terminal nodes, values within
    a few ulp of the summation-order difference."""
    io = np.load(os.path.join(GOLD, "model_tree_io.npz"))
    y1 = open(os.path.join(GOLD, "model_tree_1.yaml")).read()
    y2 = open(os.path.join(GOLD, "model_tree_2.yaml")).read()
    tree_oracle.load_model_trees(y1, y2)
    for which, name in ((0, "1"), (1, "2")):
        y, node = tree_oracle.model_tree_predict(which, io["X"])
        assert np.array_equal(node, io["node_" + name])
        assert np.allclose(y, io["y_" + name], rtol=1e-12, atol=1e-12)
    yp = open(os.path.join(GOLD, "model_tree_poly2.yaml")).read()
    tree_oracle.load_model_trees(yp, yp, degree=2)
    y, node = tree_oracle.model_tree_predict(0, io["X"])
    assert np.array_equal(node, io["node_poly2"])
    assert np.allclose(y, io["y_poly2"], rtol=1e-10, atol=1e-10)
    with pytest.raises(RuntimeError):   # model_tree.cc:50-51: 119 features against 15-parameter models
        tree_oracle.load_model_trees(y1, y2, degree=2)
        tree_oracle.model_tree_predict(0, io["X"])


def test_oracle_interaction_only_features_follow_sklearn_order(tree_oracle):
    """Polynomial_model_tree(degree 2, interaction_only) builds x_i*x_j (i<j) in PolynomialFeatures' order."""
    n = 14
    nf = n + n * (n - 1) // 2
    w = np.arange(1, nf + 2, dtype=np.float64) / 7.0
    text = "1:\n  terminal: true\n  model params: [%s]\n" % ", ".join(repr(float(v)) for v in w)
    tree_oracle.load_model_trees(text, text, degree=2, interaction_only=True)
    x = np.random.RandomState(1).uniform(-2, 2, size=(16, n))
    feats = [x] + [x[:, i:i + 1] * x[:, j:j + 1] for i in range(n) for j in range(i + 1, n)]
    want = np.hstack(feats) @ w[:-1] + w[-1]
    y, node = tree_oracle.model_tree_predict(0, x)
    assert (node == 1).all() and np.allclose(y, want, rtol=1e-12)


# ---------------------------------------------------------------------------------------------------------------------- SavedModel variables
# Writer of the TensorFlow tensor-bundle checkpoint as TensorFlow's BundleWriter lays it out (tensor_bundle.cc + lib/io/table_builder.cc:
# sorted keys, shared-prefix compression with a restart every 16 entries, uncompressed blocks with a masked-CRC32C trailer, empty metaindex
# block, index block, 48-byte footer).  FORMAT UNPINNED: restated from the published format, no TensorFlow-written file exists on this box.
def _crc32c(data):
    tbl = _crc32c.tbl
    if tbl is None:
        tbl = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ (0x82F63B78 if c & 1 else 0)
            tbl.append(c)
        _crc32c.tbl = tbl
    c = 0xFFFFFFFF
    for b in data:
        c = tbl[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


_crc32c.tbl = None


def _mask(crc):
    return ((((crc >> 15) | (crc << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def _varint(v):
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _table_block(entries, restart_interval=16):
    body, restarts, last = bytearray(), [], b""
    for i, (k, v) in enumerate(entries):
        shared = 0
        if i % restart_interval == 0:
            restarts.append(len(body))
        else:
            while shared < min(len(k), len(last)) and k[shared] == last[shared]:
                shared += 1
        body += _varint(shared) + _varint(len(k) - shared) + _varint(len(v)) + k[shared:] + v
        last = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        body += struct.pack("<I", r)
    return bytes(body + struct.pack("<I", len(restarts)))


def _entry_proto(dtype, shape, shard, offset, size, crc):
    dims = b"".join(b"\x12" + _varint(len(d)) + d for d in (b"\x08" + _varint(s) for s in shape))
    return (b"\x08" + _varint(dtype) + b"\x12" + _varint(len(dims)) + dims + (b"\x18" + _varint(shard) if shard else b"")
            + (b"\x20" + _varint(offset) if offset else b"") + b"\x28" + _varint(size) + b"\x35" + struct.pack("<I", crc))


def write_bundle(prefix, tensors, num_shards=1, entries_per_block=1000, compression=0, corrupt=None, extra_entries=()):
    """tensors: {key: (np array, dtype enum)}; tensor i goes to shard i % num_shards."""
    data = [bytearray() for _ in range(num_shards)]
    entries = [(b"", b"\x08" + _varint(num_shards) + b"\x1a\x02\x08\x01")]   # BundleHeaderProto: num_shards, version { producer: 1 }
    for i, key in enumerate(sorted(tensors)):
        arr, dtype = tensors[key]
        raw = np.ascontiguousarray(arr).tobytes()
        sh = i % num_shards
        off = len(data[sh])
        data[sh] += raw
        crc = _mask(_crc32c(raw))
        if corrupt == key:
            crc ^= 1
        entries.append((key.encode(), _entry_proto(dtype, arr.shape, sh, off, len(raw), crc)))
    entries += list(extra_entries)
    entries.sort()
    out, index = bytearray(), []

    def emit(block):
        handle = _varint(len(out)) + _varint(len(block))
        out.extend(block + bytes([compression]) + struct.pack("<I", _mask(_crc32c(block + bytes([compression])))))
        return handle

    for k in range(0, len(entries), entries_per_block):
        chunk = entries[k:k + entries_per_block]
        index.append((chunk[-1][0] + b"\x00", emit(_table_block(chunk))))     # any separator >= the block's last key
    meta = emit(_table_block([]))
    idx = emit(_table_block(index, restart_interval=1))
    foot = meta + idx
    out.extend(foot + b"\x00" * (40 - len(foot)) + struct.pack("<Q", 0xDB4775248B80FB57))
    os.makedirs(os.path.dirname(prefix), exist_ok=True)
    with open(prefix + ".index", "wb") as f:
        f.write(out)
    for s in range(num_shards):
        with open(prefix + ".data-%05d-of-%05d" % (s, num_shards), "wb") as f:
            f.write(data[s])


def _actor_tensors(seed=3, dims=(17, 512, 512, 2)):
    rng = np.random.default_rng(seed)
    t, blob = {}, []
    for l in range(len(dims) - 1):
        w = rng.standard_normal((dims[l], dims[l + 1])).astype(np.float32)
        b = rng.standard_normal(dims[l + 1]).astype(np.float32)
        t["layer_with_weights-%d/kernel/.ATTRIBUTES/VARIABLE_VALUE" % l] = (w, 1)
        t["layer_with_weights-%d/bias/.ATTRIBUTES/VARIABLE_VALUE" % l] = (b, 1)
        blob += [w.ravel(), b.ravel()]
    t["_CHECKPOINTABLE_OBJECT_GRAPH"] = (np.frombuffer(b"\x0a\x10object-graph-pb", dtype=np.uint8), 7)   # DT_STRING payload, skipped
    return t, np.concatenate(blob)


def test_crc32c_known_answers():
    # RFC 3720 B.4 test vectors of CRC32C
    assert _crc32c(b"123456789") == 0xE3069283
    assert _crc32c(bytes(32)) == 0x8A9136AA
    assert _crc32c(bytes([0xFF] * 32)) == 0x62A8AB43


@pytest.mark.parametrize("num_shards,entries_per_block", [(1, 1000), (2, 3), (1, 1)])
def test_saved_model_actor_round_trip(tmp_path, num_shards, entries_per_block):
    from tf_ode_b200 import read_saved_model_actor
    t, blob = _actor_tensors()
    write_bundle(str(tmp_path / "actor" / "variables" / "variables"), t, num_shards, entries_per_block)
    w = read_saved_model_actor(str(tmp_path / "actor"))
    assert w.dtype == np.float32 and w.size == 272898
    assert np.array_equal(w.view(np.uint32), blob.view(np.uint32))


def test_saved_model_errors_are_loud(tmp_path):
    from tf_ode_b200 import read_saved_model_actor
    t, _ = _actor_tensors()
    d = tmp_path / "a"
    prefix = str(d / "variables" / "variables")
    with pytest.raises(TfodeError, match="Can't open"):
        read_saved_model_actor(str(d))
    write_bundle(prefix, t, corrupt="layer_with_weights-1/kernel/.ATTRIBUTES/VARIABLE_VALUE")
    with pytest.raises(TfodeError, match="CRC32C mismatch"):
        read_saved_model_actor(str(d))
    write_bundle(prefix, t, compression=1)
    with pytest.raises(TfodeError, match="compressed block"):
        read_saved_model_actor(str(d))
    missing = {k: v for k, v in t.items() if "weights-2/bias" not in k}
    write_bundle(prefix, missing)
    with pytest.raises(TfodeError, match="no variable 'layer_with_weights-2/bias"):
        read_saved_model_actor(str(d))
    t64 = dict(t)
    k0 = "layer_with_weights-0/kernel/.ATTRIBUTES/VARIABLE_VALUE"
    t64[k0] = (t[k0][0].astype(np.float64), 2)
    write_bundle(prefix, t64)
    with pytest.raises(TfodeError, match="not DT_FLOAT"):
        read_saved_model_actor(str(d))
    tc, _ = _actor_tensors(dims=(19, 512, 512, 1))     # the critic of rover_training_1.py:41-52 saved in the actor's place
    write_bundle(prefix, tc)
    with pytest.raises(TfodeError, match=r"shape \[19,512,\] instead of \[17,512,\]"):
        read_saved_model_actor(str(d))
    write_bundle(prefix, t)
    raw = bytearray(open(prefix + ".index", "rb").read())
    raw[-1] ^= 0xFF
    open(prefix + ".index", "wb").write(raw)
    with pytest.raises(TfodeError, match="bad table magic"):
        read_saved_model_actor(str(d))
    write_bundle(prefix, t)
    raw = bytearray(open(prefix + ".index", "rb").read())
    raw[10] ^= 0x40
    open(prefix + ".index", "wb").write(raw)
    with pytest.raises(TfodeError, match="checksum mismatch"):
        read_saved_model_actor(str(d))
    write_bundle(prefix, t)
    with open(prefix + ".data-00000-of-00001", "r+b") as f:
        f.truncate(1000)
    with pytest.raises(TfodeError, match="outside its data shard"):
        read_saved_model_actor(str(d))


# ----------------------------------------------------------------------------------------------------------------------
# PRINT_TRANSITIONS / PRINT_STATE_AND_ACTIONS row formats (src/rover_1_tf.cc:125-136,190-195), read by scripts/bin/monitor-policies
# ----------------------------------------------------------------------------------------------------------------------
def test_print_transitions_row_formats_match_the_reference_printf_calls():
    from tf_ode_b200 import abi, format_transitions
    rs = np.random.RandomState(4)
    rows = np.zeros((5, abi.TRANSITION_WIDTH), np.float32)
    rows[:, 0] = np.arange(5)
    rows[:, 1:18] = rs.randn(5, 17) * 40
    rows[:, 18:20] = rs.uniform(-1, 1, (5, 2)) * np.array([15, 25])
    rows[:, 20] = rs.randn(5); rows[:, 21] = (0, 0, 1, 0, 1)
    rows[:, 22:39] = rs.randn(5, 17) * 40
    rows[2, 1] = -0.0000004; rows[3, 22] = 123456.789
    # the reference: printf("%f ", s_i) x17, printf("%f %f", steer, torque), printf(" %f", s'_i) x17, printf("\n")
    want = "".join("".join("%f " % float(v) for v in r[1:18]) + "%f %f" % (float(r[18]), float(r[19])) + "".join(" %f" % float(v) for v in r[22:39]) + "\n" for r in rows)
    assert format_transitions(rows, 0) == want
    # PRINT_STATE_AND_ACTIONS: printf("%f ", s_i) x17, printf("%f %f\n", steer, torque)
    want = "".join("".join("%f " % float(v) for v in r[1:18]) + "%f %f\n" % (float(r[18]), float(r[19])) for r in rows)
    assert format_transitions(rows, 1) == want
    assert len(format_transitions(rows[:1], 0).split()) == 36 and len(format_transitions(rows[:1], 1).split()) == 19
    with pytest.raises(Exception):
        format_transitions(rows, 7)
