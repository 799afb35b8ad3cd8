"""Host side above the device ABI (ipk_b200/host): the raxml-ng posterior loader against the REFERENCE's own csv
reader (oracle/_ref/ref_csv_*, built from /root/reference's csv.h as ar.cpp instantiates it) and against committed
fixtures of it; the .ipk writer/reader round trip; and -- on a GPU -- the ipk-b200 executable end to end."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
REF_CSV = {4: os.path.join(ROOT, "oracle", "_ref", "ref_csv_dna"), 20: os.path.join(ROOT, "oracle", "_ref", "ref_csv_aa")}
AA_RAXML = "ARNDCQEGHILKMFPSTWYV"


def write_tsv(path, probs, labels, sigma, rng=None, tricky=False):
    """probs [n, W, sigma] in the FILE's column order; numbers printed the way raxml-ng does (--precision 9), plus a few
    other spellings the csv parser accepts when `tricky`"""
    names = "ACGT" if sigma == 4 else AA_RAXML
    with open(path, "w") as f:
        f.write("Node\tSite\tState\t" + "\t".join("p_" + c for c in names) + "\n")
        for n, lab in enumerate(labels):
            for s in range(probs.shape[1]):
                row = probs[n, s]
                cells = []
                for j, p in enumerate(row):
                    t = f"{p:.9f}"
                    if tricky and rng is not None:
                        r = rng.integers(0, 12)
                        if r == 0: t = f"{p:.6e}"
                        elif r == 1: t = f" {p:.5f} "
                        elif r == 2: t = f"{p:.17f}"
                        elif r == 3 and p < 1e-3: t = f"{p:.3E}"
                        elif r == 4: t = "+" + t
                    cells.append(t)
                f.write(f"{lab}\t{s + 1}\t{names[int(np.argmax(row))]}\t" + "\t".join(cells) + "\n")
            if tricky and n == 1:
                f.write("\n.a comment line as single_and_empty_line_comment<'.'> skips it\n")


def random_probs(rng, n, W, sigma):
    p = rng.dirichlet(np.full(sigma, 0.3), size=(n, W))
    p[rng.random((n, W)) < 0.1] = np.eye(sigma)[0]                       # exact 1 / 0 columns -> log10(0) = -inf
    return p


def ref_parse(tsv, sigma, tmp_path):
    out = str(tmp_path / "ref.f32")
    subprocess.run([REF_CSV[sigma], tsv, out], check=True)
    return np.fromfile(out, np.float32)


@pytest.mark.parametrize("sigma", [4, 20])
def test_loader_matches_reference_csv_reader(sigma, tmp_path):
    if not os.path.exists(REF_CSV[sigma]):
        pytest.skip("oracle/_ref/ref_csv_* not built (needs /root/reference)")
    from ipk_b200 import hostlib
    rng = np.random.default_rng(3 + sigma)
    labels = [f"Node{i}" for i in range(1, 8)]
    probs = random_probs(rng, len(labels), 37, sigma)
    tsv = str(tmp_path / "probs.tsv")
    write_tsv(tsv, probs, labels, sigma, rng, tricky=True)
    want = ref_parse(tsv, sigma, tmp_path).reshape(len(labels), 37, sigma)
    got_labels, got = hostlib.read_ar(tsv, sigma)
    assert got_labels == labels
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), "loader differs bitwise from the reference's csv reader"
    # a subset, in another order, as db_builder asks for nodes
    pick = ["Node5", "Node2", "Node7"]
    _, sub = hostlib.read_ar(tsv, sigma, pick, threads=2)
    assert np.array_equal(sub.view(np.uint32), want[[4, 1, 6]].view(np.uint32))


@pytest.mark.parametrize("sigma", [4, 20])
def test_loader_matches_committed_fixture(sigma):
    """the same comparison against outputs of the reference's reader committed under tests/golden (make_golden_ar.py)"""
    from ipk_b200 import hostlib
    tsv = os.path.join(GOLD, f"ar_{'dna' if sigma == 4 else 'aa'}.tsv")
    want = np.fromfile(os.path.join(GOLD, f"ar_{'dna' if sigma == 4 else 'aa'}.f32"), np.float32)
    labels, got = hostlib.read_ar(tsv, sigma)
    assert len(labels) == 3
    assert np.array_equal(got.reshape(-1).view(np.uint32), want.view(np.uint32))


def test_parse_float_known_answers():
    """csv.h:1024-1085 accumulates in float: not the correctly rounded value for long digit strings"""
    from ipk_b200 import hostlib
    f = np.float32
    assert hostlib.parse_float("0.5") == 0.5 and hostlib.parse_float("1") == 1.0 and hostlib.parse_float("-2.25") == -2.25
    x = f(0)
    pos = f(1)
    for ch in "123456789":                    # 0.123456789 the reference's way
        pos = f(pos / f(10))
        x = f(x + f(int(ch) * pos))
    assert hostlib.parse_float("0.123456789") == float(x)
    # 1e-3: base 0.1f, e = 3 -> x *= base (e = 2); base = base*base (e = 1); x *= base   (csv.h:1058-1070)
    b = f(0.1)
    assert hostlib.parse_float("1e-3") == float(f(f(f(1) * b) * f(b * b)))
    with pytest.raises(hostlib.HostError):
        hostlib.parse_float("0.5x")
    with pytest.raises(hostlib.HostError):
        hostlib.parse_float("abc")


def test_loader_errors(tmp_path):
    from ipk_b200 import hostlib
    with pytest.raises(hostlib.HostError):
        hostlib.read_ar(str(tmp_path / "missing.tsv"), 4)
    tsv = str(tmp_path / "p.tsv")
    write_tsv(tsv, random_probs(np.random.default_rng(1), 2, 5, 4), ["A1", "A2"], 4)
    with pytest.raises(hostlib.HostError, match="Could not read the AR matrix"):
        hostlib.read_ar(tsv, 4, ["A1", "nope"])
    with open(tsv, "a") as f:
        f.write("A3\t1\tA\t0.5\t0.5\n")       # too few columns
    with pytest.raises(hostlib.HostError):
        hostlib.read_ar(tsv, 4, ["A3"])


def test_ipk_round_trip(tmp_path):
    from ipk_b200 import hostlib
    rng = np.random.default_rng(9)
    nk = 1000
    cnt = rng.integers(1, 9, nk)
    offsets = np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64)
    ne = int(offsets[-1])
    keys = rng.permutation(1 << 16)[:nk].astype(np.uint64)
    fv = np.sort(rng.random(nk).astype(np.float32))
    branch = rng.integers(0, 500, ne).astype(np.uint32)
    score = (-5 * rng.random(ne)).astype(np.float32)
    path = str(tmp_path / "DB.ipk")
    hostlib.write_ipk(path, keys, fv, offsets, branch, score, 8, 1.5, "DNA", "((a:1,b:2):0.5,c:3);", [1, 1, 3, 1, 5], [1, 2, 3.5, 3, 6.5])
    d = hostlib.read_ipk(path)
    assert d["sequence_type"] == "DNA" and d["tree"] == "((a:1,b:2):0.5,c:3);" and d["kmer_size"] == 8 and d["omega"] == 1.5
    assert list(d["index_num_nodes"]) == [1, 1, 3, 1, 5] and list(d["index_len"]) == [1, 2, 3.5, 3, 6.5]
    for name, a in (("keys", keys), ("fv", fv), ("offsets", offsets), ("branch", branch), ("score", score)):
        assert np.array_equal(d[name], a), name
    assert os.path.getsize(path) == 8 + 22 + 2 + 4 + 4 + (8 + 3) + 8 + 5 * 16 + (8 + 20) + 8 + 4 + 8 + 8 + nk * 20 + ne * 8
    with open(path, "r+b") as f:              # truncated file -> format error, not a crash
        f.truncate(os.path.getsize(path) - 5)
    with pytest.raises(hostlib.HostError):
        hostlib.read_ipk(path)
    hostlib.write_ipk(path, [], [], [0], [], [], 8, 1.5)
    assert len(hostlib.read_ipk(path)["keys"]) == 0
    # the preamble of a boost binary_oarchive with default flags (db_builder.cpp:176-177)
    raw = open(path, "rb").read(40)
    assert raw[:8] == (22).to_bytes(8, "little") and raw[8:30] == b"serialization::archive" and raw[30:32] == (19).to_bytes(2, "little")
    assert raw[32:36] == bytes([4, 8, 4, 8]) and raw[36:40] == (1).to_bytes(4, "little")


def test_db_tools_diff_and_dump(tmp_path):
    """ipk-b200-db: the checks of tools/src/diff.cpp:24-295 (+ BASELINE.json's boundary rule) and the listing of dump.cpp:18-33"""
    from ipk_b200 import hostlib
    tool = os.path.join(os.path.dirname(hostlib.EXE_PATH), "ipk-b200-db")
    tree = "((a:1,b:2):0.5,c:3);"                    # post-order a0 b1 (ab)2 c3 root4; pre-order root0 (ab)1 a2 b3 c4
    k, omega = 3, 1.5
    thr = float(np.log10(np.float32((omega / 4) ** k)))
    keys = np.array([5, 9, 27], np.uint64)          # AAC.. in base 4: 5 = 0,1,1 -> "ACC"; 9 = "AGC"; 27 = "CGT"
    offsets = np.array([0, 2, 3, 5], np.uint64)
    branch = np.array([0, 3, 1, 2, 0], np.uint32)
    score = np.array([-0.5, -1.0, -0.25, thr + 2e-6, -0.75], np.float32)
    fv = np.array([0.1, 0.2, 0.3], np.float32)
    pa, pb, pc = (str(tmp_path / n) for n in ("A.ipk", "B.ipk", "C.ipk"))
    hostlib.write_ipk(pa, keys, fv, offsets, branch, score, k, omega, "DNA", tree, [1, 1, 3, 1, 5], [0, 0, 3, 0, 6.5])
    # B: same database, k-mers in another order, one score off by 1e-3 (inside the reference's EPS), the boundary entry missing
    order = [2, 0, 1]
    cnt = np.diff(offsets.astype(np.int64))
    sel = np.concatenate([np.arange(offsets[q], offsets[q + 1]) for q in order]).astype(np.int64)
    b_branch, b_score = branch[sel].copy(), score[sel].copy()
    b_score[np.where((b_branch == 3))[0][0]] += np.float32(1e-3)
    drop = int(np.where(b_score == np.float32(thr + 2e-6))[0][0])
    b_cnt = cnt[order].copy(); b_cnt[0] -= 1
    hostlib.write_ipk(pb, keys[order], fv[order], np.concatenate(([0], np.cumsum(b_cnt))), np.delete(b_branch, drop), np.delete(b_score, drop),
                      k, omega, "DNA", tree, [1, 1, 3, 1, 5], [0, 0, 3, 0, 6.5])
    r = subprocess.run([tool, "diff", pa, pa], capture_output=True, text=True)
    assert r.returncode == 0 and "Phylo-k-mer scores:\tOK\t0" in r.stdout and "Sequence type:\tOK\tDNA\tDNA" in r.stdout
    r = subprocess.run([tool, "diff", pa, pb, "-v"], capture_output=True, text=True)
    assert r.returncode == 2 and "Number of phylo-k-mers:\tDIFF\t5\t4" in r.stdout and "Phylo-k-mer scores:\tDIFF\t1" in r.stdout
    assert "\t27\tCGT\t2\t" in r.stdout                 # the entry B lacks, decoded
    r = subprocess.run([tool, "diff", pa, pb, "--boundary", "1e-5"], capture_output=True, text=True)
    assert "Phylo-k-mer scores:\tOK\t0" in r.stdout and "Boundary cases (a score within 1e-05 of the threshold):\t1" in r.stdout
    r = subprocess.run([tool, "diff", pa, pb, "--boundary", "1e-5", "--eps", "0"], capture_output=True, text=True)
    assert "Phylo-k-mer scores:\tDIFF\t1" in r.stdout      # bit equality: the 1e-3 shift shows
    r = subprocess.run([tool, "dump", pa], capture_output=True, text=True)
    assert r.returncode == 0
    lines = r.stdout.splitlines()
    assert lines[0] == "ACC" and lines[1].split("\t")[1:] == [f"{10 ** -0.5:g}", "2"] and lines[2].split("\t")[2] == "4"      # branches as pre-order ids
    assert lines[3] == "AGC" and lines[5] == "CGT"
    r = subprocess.run([tool, "diff", pa, str(tmp_path / "missing.ipk")], capture_output=True, text=True)
    assert r.returncode == 1 and "Exception" in r.stderr


def test_ipk_round_trip_with_positions(tmp_path):
    """KEEP_POSITIONS databases (ipk-aa-pos): entries are {branch, score, position}; the reader tells the two layouts apart by
    the length of the file"""
    from ipk_b200 import hostlib
    rng = np.random.default_rng(19)
    nk = 3000
    cnt = rng.integers(1, 7, nk)
    offsets = np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64)
    ne = int(offsets[-1])
    keys = rng.permutation(1 << 15)[:nk].astype(np.uint64)
    fv = np.sort(rng.random(nk).astype(np.float32))
    branch = rng.integers(0, 300, ne).astype(np.uint32)
    score = (-5 * rng.random(ne)).astype(np.float32)
    pos = rng.integers(0, 800, ne).astype(np.uint32)
    a, b = str(tmp_path / "pos.ipk"), str(tmp_path / "plain.ipk")
    hostlib.write_ipk(a, keys, fv, offsets, branch, score, 6, 1.5, "Proteins", "(a:1,b:2);", position=pos, threads=3)
    hostlib.write_ipk(b, keys, fv, offsets, branch, score, 6, 1.5, "Proteins", "(a:1,b:2);")
    assert os.path.getsize(a) - os.path.getsize(b) == 2 * ne
    d = hostlib.read_ipk(a)
    assert np.array_equal(d["position"], pos) and np.array_equal(d["branch"], branch) and np.array_equal(d["score"], score)
    assert hostlib.read_ipk(b)["position"] is None


def test_ipk_writer_threads_agree(tmp_path):
    """the multi-threaded writer (runs of k-mers at prefix-summed offsets) produces the bytes of the single-threaded one"""
    from ipk_b200 import hostlib
    rng = np.random.default_rng(11)
    nk = 50000
    cnt = rng.integers(0, 40, nk)
    cnt[rng.integers(0, nk, 20)] = 5000                 # a few long k-mers straddle block and thread boundaries
    offsets = np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64)
    ne = int(offsets[-1])
    keys = rng.permutation(1 << 20)[:nk].astype(np.uint64)
    fv = np.sort(rng.random(nk).astype(np.float32))
    branch = rng.integers(0, 60000, ne).astype(np.uint32)
    score = (-5 * rng.random(ne)).astype(np.float32)
    files = []
    for t in (1, 3, 16):
        path = str(tmp_path / f"DB{t}.ipk")
        hostlib.write_ipk(path, keys, fv, offsets, branch, score, 10, 1.5, "DNA", "(a:1,b:2);", [1, 1, 3], [0, 0, 3], threads=t)
        files.append(open(path, "rb").read())
    assert files[0] == files[1] == files[2]
    d = hostlib.read_ipk(str(tmp_path / "DB16.ipk"))
    assert np.array_equal(d["branch"], branch) and np.array_equal(d["score"], score) and np.array_equal(d["offsets"], offsets)


# ---- host orchestration: tree -> ghost nodes -> groups -> AR labels -> tree index (ipkhost.h build plan) ---------------
FIVE = "((A:1,B:2)n1:0.5,(C:3,(D:4,E:5)n3:0.25)n2:0.75)root;"


def test_plan_five_taxa_by_hand():
    """extended_tree.cpp:76-155 + db_builder.cpp:495-553 + :191-197 on a tree small enough to do on paper.
    Original post-order: A0 B1 n1=2 C3 D4 E5 n3=6 n2=7 root8; ghost names count from node_count + 1 = 10 in that order."""
    from ipk_b200 import hostlib
    p = hostlib.plan(FIVE)
    assert p["node_count"] == 9 and p["n_groups"] == 8
    # post-order of the EXTENDED tree: X1 of a node comes before the ghosts of its descendants, X0 after them
    assert p["ext_labels"] == ["19_X1", "18_X0", "11_X1", "10_X0", "15_X1", "14_X0", "39_X1", "38_X0", "23_X1", "22_X0",
                               "35_X1", "34_X0", "27_X1", "26_X0", "31_X1", "30_X0"]
    assert list(p["node_branch"]) == [2, 2, 0, 0, 1, 1, 7, 7, 3, 3, 6, 6, 4, 4, 5, 5]
    assert p["ar_labels"] == p["ext_labels"]              # no AR tree given
    assert list(p["index_num_nodes"]) == [1, 1, 3, 1, 1, 1, 3, 5, 9]
    assert list(p["index_len"]) == [0, 0, 3, 0, 0, 0, 9, 12.25, 16.5]
    assert p["newick"] == FIVE.replace(":0.5,", ":0.5,")       # round trip of labels, lengths and child order
    # ghost branch lengths (extended_tree.cpp:35-73): leaf A (1): x0 = x1 = 0.5; n1 (0.5, two leaves, 3 below): x0 0.25, x1 (3 + 0.25 * 2) / 2
    ext = p["extended_newick"]
    for piece in ("11_X1:0.5", "A:0.5)10_X0:0.5", "19_X1:1.75", "18_X0:0.25", "12_X2:0.01,13_X3:0.01"):
        assert piece in ext, (piece, ext)
    inner, outer = hostlib.plan(FIVE, ghosts="inner-only"), hostlib.plan(FIVE, ghosts="outer-only")
    assert list(inner["node_branch"]) == [0, 1, 2, 3, 4, 5, 6, 7] and all(l.endswith("_X0") for l in inner["ext_labels"])
    assert list(outer["node_branch"]) == [2, 0, 1, 7, 3, 6, 4, 5] and all(l.endswith("_X1") for l in outer["ext_labels"])


def _relabel(newick, fn):
    """every label of a newick string through fn (labels here contain no quotes)"""
    import re
    return re.sub(r"([(),])([^(),:;]+)", lambda m: m.group(1) + fn(m.group(2)), newick)


def _top_children(newick):
    """texts of the root's children"""
    body = newick[1:newick.rindex(")")]
    out, depth, start = [], 0, 0
    for i, c in enumerate(body):
        depth += c == "("
        depth -= c == ")"
        if c == "," and depth == 0:
            out.append(body[start:i]); start = i + 1
    out.append(body[start:])
    return out


def test_plan_maps_ar_labels_and_reroots():
    """ar::map_nodes (ar.cpp:790-834) over a relabelled copy of the extended tree; and the same tree as raxml-ng returns it,
    unrooted ((X, Y)root -> (Y, x1, x2), main.cpp:166-178), which reroot_tree (extended_tree.cpp:186-205) restores"""
    from ipk_b200 import hostlib
    ext = hostlib.plan(FIVE)["extended_newick"]
    ar_rooted = _relabel(ext, lambda l: "AR" + l)
    p = hostlib.plan(FIVE, ar_rooted)
    assert p["ar_labels"] == ["AR" + l for l in p["ext_labels"]]
    x, y = _top_children(ar_rooted)                     # x = (...)AR18_X0:len
    x_kids = _top_children(x[: x.rindex(")") + 1])
    unrooted = "(" + ",".join([y] + x_kids) + ")ARroot;"
    q = hostlib.plan(FIVE, unrooted)
    want = ["AR" + l for l in p["ext_labels"]]
    want[p["ext_labels"].index("18_X0")] = "ARroot"     # the dropped node's place is taken by the old root, the new one is "added_root"
    assert q["ar_labels"] == want
    with pytest.raises(hostlib.HostError, match="differ in the number of nodes"):
        hostlib.plan(FIVE, "(a,b);")


def test_plan_rejects_unrooted_reference_tree():
    from ipk_b200 import hostlib
    with pytest.raises(hostlib.HostError, match="not rooted"):
        hostlib.plan("(A:1,B:1,C:1);")
    assert hostlib.plan("(A:1,B:1,C:1);", use_unrooted=True)["n_groups"] == 3
    with pytest.raises(hostlib.HostError, match="newick"):
        hostlib.plan("((A:1,B:1);")


def _ipk_py_argv(exe, tmp_path, ar_dir, extra=()):
    """the command ipk.py:292-329 builds"""
    return [exe, "--ar-binary", "raxml-ng", "--refalign", str(tmp_path / "ref.fasta"), "-t", str(tmp_path / "tree.newick"), "-w", str(tmp_path),
            "-k", "6", "--model", "GTR", "--alpha", "1.0", "--categories", "4", "--reduction-ratio", "0.99", "--omega", "1.5", "--mif0", "--both",
            "-u", "1.0", "-j", "2", "-o", str(tmp_path / "DB.ipk"), "-v", "1", "--ar-dir", ar_dir, *extra]


def test_executable_accepts_the_command_line_ipk_py_builds(tmp_path):
    """CPU: every option of ipk/src/command_line.cpp:83-147 parses; without a GPU the run ends at the device, not at the argv"""
    from ipk_b200 import hostlib
    exe = os.path.join(os.path.dirname(hostlib.EXE_PATH), "ipk-dna")
    assert os.path.exists(exe)
    (tmp_path / "tree.newick").write_text(FIVE + "\n")
    ar_dir = tmp_path / "AR"
    ar_dir.mkdir()
    p = hostlib.plan(FIVE)
    rng = np.random.default_rng(2)
    write_tsv(str(ar_dir / "x.raxml.ancestralProbs"), random_probs(rng, len(p["ext_labels"]), 30, 4), p["ext_labels"], 4)
    (ar_dir / "x.raxml.ancestralTree").write_text(p["extended_newick"] + "\n")
    extra = ["--ar-parameters", '"--blopt', 'nr_safe"', "--no-reduction", "--use-unrooted", "--uncompressed", "--on-disk", "--DCLA"]
    r = subprocess.run(_ipk_py_argv(exe, tmp_path, str(ar_dir), extra), capture_output=True, text=True)
    assert "unknown option" not in r.stderr and "usage:" not in r.stderr, r.stderr
    assert "16 ghost nodes in 8 branch groups of a 9-node tree" in r.stderr, r.stderr
    assert (tmp_path / "extended_trees" / "extended_tree.newick").read_text().strip() == p["extended_newick"]
    import torch
    if not torch.cuda.is_available():
        assert r.returncode == 1 and "Exception" in r.stderr            # no CPU fallback: the device is where it stops
    # --ar-only ends after locating the reconstruction; a missing --ar-dir and unsupported modes are errors with exit status 1
    r = subprocess.run(_ipk_py_argv(exe, tmp_path, str(ar_dir), ["--ar-only"]), capture_output=True, text=True)
    assert r.returncode == 0 and "--ar-only requested" in r.stderr
    r = subprocess.run(_ipk_py_argv(exe, tmp_path, str(tmp_path / "nope")), capture_output=True, text=True)
    assert r.returncode == 1 and "No such directory" in r.stderr
    r = subprocess.run(_ipk_py_argv(exe, tmp_path, str(ar_dir), ["--merge-branches"]), capture_output=True, text=True)
    assert r.returncode == 1 and "only supported for IPK compiled with the KEEP_POSITIONS flag" in r.stderr      # main.cpp:31-37
    r = subprocess.run([exe, "--refalign", "a", "-t", str(tmp_path / "tree.newick")], capture_output=True, text=True)
    assert r.returncode == 1 and "--ar-dir is required" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("sigma,bits,k,omega,filt", [(4, 2, 8, 1.5, "--mif0"), (20, 5, 3, 1.5, "--random")])
def test_ipk_b200_executable_end_to_end(sigma, bits, k, omega, filt, tmp_path):
    """AR file + node map -> ipk-b200 -> DB.ipk, compared entry for entry with the oracle build on the loader's matrices"""
    import oraclelib as O
    from parity import assert_db_equal
    from types import SimpleNamespace
    from ipk_b200 import hostlib, capi, synth
    n_taxa, W = 6, 60
    logp, nb = synth.posteriors(n_taxa, W, sigma, seed=5)
    probs = np.power(10.0, logp.astype(np.float64))
    if sigma == 20:                            # the file is in raxml-ng's column order; logp is in IPK's
        order = "RHKDESTNQCGPAILMFWYV"
        probs = probs[:, :, [order.index(c) for c in AA_RAXML]]
    labels = [f"Node{i + 1}" for i in range(len(nb))]
    tsv, nmap, out = str(tmp_path / "ar.raxml.ancestralProbs"), str(tmp_path / "node_map.tsv"), str(tmp_path / "DB.ipk")
    write_tsv(tsv, probs, labels, sigma)
    with open(nmap, "w") as f:
        f.write("# AR node label <TAB> original branch (post-order id)\n")
        for lab, b in zip(labels, nb):
            f.write(f"{lab}\t{int(b)}\n")
    r = subprocess.run([hostlib.EXE_PATH, "--ar-probs", tsv, "--node-map", nmap, "--states", "dna" if sigma == 4 else "aa", "-k", str(k),
                        "--omega", str(omega), filt, "-o", out, "--node-count", str(2 * n_taxa - 1)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    d = hostlib.read_ipk(out)
    got = SimpleNamespace(keys=d["keys"], fv=d["fv"], offsets=d["offsets"], branch=d["branch"], score=d["score"],
                          n_kmers=len(d["keys"]), n_entries=len(d["branch"]))
    _, parsed = hostlib.read_ar(tsv, sigma, labels)
    thr, lthr = capi.score_threshold(omega, sigma, k)
    want = O.build(parsed, nb, W, sigma, bits, k, lthr, thr, 2 * n_taxa - 1, 0 if filt == "--mif0" else 1, 2)
    assert want.n_entries > 100
    assert_db_equal(got, want)
    assert d["kmer_size"] == k and d["sequence_type"] == ("DNA" if sigma == 4 else "Proteins")
    # error convention of the reference binary: message on stderr, exit status 1 (main.cpp:225-246)
    r = subprocess.run([hostlib.EXE_PATH, "--ar-probs", tsv + ".missing", "--node-map", nmap], capture_output=True, text=True)
    assert r.returncode == 1 and "Exception" in r.stderr


def _synth_newick(n_taxa, seed=5):
    """the generator's tree (post-order ids = branch ids of synth.posteriors) as newick; children in ascending id order keep the ids"""
    from ipk_b200 import synth
    parent, blen = synth.tree(n_taxa, seed)
    kids = {}
    for v, p in enumerate(parent):
        if p != 0xffffffff:
            kids.setdefault(int(p), []).append(v)
    def text(v):
        if v not in kids:
            return f"T{v}:{float(blen[v]):.6g}"
        inner = ",".join(text(c) for c in sorted(kids[v]))
        return f"({inner})" + ("" if parent[v] == 0xffffffff else f":{float(blen[v]):.6g}")
    return text(len(parent) - 1) + ";"


@pytest.mark.gpu
@pytest.mark.parametrize("ghosts", ["both", "outer-only"])
def test_reference_command_line_end_to_end(ghosts, tmp_path):
    """tree + raxml-ng outputs -> `ipk-dna <the argv of ipk.py:292-329>` -> DB.ipk: ghost injection, grouping, AR label mapping,
    posteriors, the device path and the header's tree index, against the oracle build over the same plan"""
    import oraclelib as O
    from parity import assert_db_equal
    from types import SimpleNamespace
    from ipk_b200 import hostlib, capi, synth
    n_taxa, W, sigma, k, omega = 7, 50, 4, 6, 1.5
    newick = _synth_newick(n_taxa)
    (tmp_path / "tree.newick").write_text(newick + "\n")
    plan0 = hostlib.plan(newick, ghosts=ghosts)
    assert plan0["node_count"] == 2 * n_taxa - 1
    logp, nb = synth.posteriors(n_taxa, W, sigma, seed=5)               # rows 2b (X0), 2b + 1 (X1) of branch b
    all_ext = hostlib.plan(newick)["ext_labels"]
    all_branch = hostlib.plan(newick)["node_branch"]
    row_of = {l: 2 * int(b) + (1 if l.endswith("_X1") else 0) for l, b in zip(all_ext, all_branch)}
    ar_dir = tmp_path / "AR"
    ar_dir.mkdir()
    file_labels = sorted(row_of, key=row_of.get)                         # the file lists the nodes in ITS order, not the build's
    write_tsv(str(ar_dir / "ext.phylip.raxml.ancestralProbs"), np.power(10.0, logp.astype(np.float64))[[row_of[l] for l in file_labels]],
              ["AR" + l for l in file_labels], sigma)
    (ar_dir / "ext.phylip.raxml.ancestralTree").write_text(_relabel(hostlib.plan(newick)["extended_newick"], lambda l: "AR" + l) + "\n")
    exe = os.path.join(os.path.dirname(hostlib.EXE_PATH), "ipk-dna")
    argv = _ipk_py_argv(exe, tmp_path, str(ar_dir))
    argv[argv.index("--both")] = "--" + ghosts
    r = subprocess.run(argv, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    d = hostlib.read_ipk(str(tmp_path / "DB.ipk"))
    assert d["tree"] == plan0["newick"] and d["kmer_size"] == 6 and d["sequence_type"] == "DNA"
    assert np.array_equal(d["index_num_nodes"], plan0["index_num_nodes"]) and np.array_equal(d["index_len"], plan0["index_len"])
    assert len(d["index_num_nodes"]) == 2 * n_taxa - 1 and d["index_num_nodes"][-1] == 2 * n_taxa - 1
    _, parsed = hostlib.read_ar(str(ar_dir / "ext.phylip.raxml.ancestralProbs"), sigma, ["AR" + l for l in plan0["ext_labels"]])
    thr, lthr = capi.score_threshold(omega, sigma, k)
    want = O.build(parsed, plan0["node_branch"], W, sigma, 2, k, lthr, thr, 2 * n_taxa - 1, 0, 2)
    got = SimpleNamespace(keys=d["keys"], fv=d["fv"], offsets=d["offsets"], branch=d["branch"], score=d["score"],
                          n_kmers=len(d["keys"]), n_entries=len(d["branch"]))
    assert want.n_entries > 100
    assert_db_equal(got, want)
    # the multi-GPU build of the same executable (ranks = threads, one per listed device; here three on the one device):
    # key-range shards merged by (filter value, key) while serializing -- the same file, byte for byte
    argv3 = argv + ["--devices", "0,0,0"]
    argv3[argv3.index("-o") + 1] = str(tmp_path / "DB3.ipk")
    r = subprocess.run(argv3, capture_output=True, text=True)
    assert r.returncode == 0 and "on 3 devices" in r.stderr, r.stderr
    assert open(tmp_path / "DB3.ipk", "rb").read() == open(tmp_path / "DB.ipk", "rb").read()
    # ... and over distinct devices where the box has them (direct peer access between the ranks' buffers)
    import torch
    n_dev = torch.cuda.device_count()
    if n_dev >= 2:
        devs = ",".join(str(i) for i in range(min(n_dev, 4)))
        argvn = argv + ["--devices", devs]
        argvn[argvn.index("-o") + 1] = str(tmp_path / "DBn.ipk")
        r = subprocess.run(argvn, capture_output=True, text=True)
        assert r.returncode == 0 and f"on {min(n_dev, 4)} devices" in r.stderr, r.stderr
        assert open(tmp_path / "DBn.ipk", "rb").read() == open(tmp_path / "DB.ipk", "rb").read()


def _random_shards(rng, n_shards, n_kmers_each, with_ties=True):
    """disjoint key sets, each shard in ascending (fv, key); a few exactly equal filter values across and inside shards"""
    from types import SimpleNamespace
    keys_all = rng.permutation(1 << 20)[: n_shards * n_kmers_each].astype(np.uint64)
    shards = []
    for s in range(n_shards):
        keys = keys_all[s * n_kmers_each:(s + 1) * n_kmers_each]
        fv = rng.normal(size=n_kmers_each).astype(np.float32)
        if with_ties:
            fv[rng.integers(0, n_kmers_each, n_kmers_each // 10)] = np.float32(0.125)
            fv[:3] = [-0.0, 0.0, -1.5]
        img = fv.view(np.uint32).astype(np.int64)
        o = np.lexsort((keys, np.where(img >> 31, img ^ 0xffffffff, img ^ 0x80000000)))     # the order the device leaves a shard in
        keys, fv = keys[o], fv[o]
        cnt = rng.integers(1, 6, n_kmers_each)
        offsets = np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64)
        ne = int(offsets[-1])
        shards.append(SimpleNamespace(keys=keys, fv=fv, offsets=offsets, branch=rng.integers(0, 900, ne).astype(np.uint32),
                                      score=(-4 * rng.random(ne)).astype(np.float32)))
    return shards


@pytest.mark.parametrize("n_shards,threads", [(1, 1), (2, 1), (3, 4), (8, 16)])
def test_merge_order_is_the_sort_by_filter_value_and_key(n_shards, threads, tmp_path):
    """ipkhost_merge_order (the N-way merge of merge_stage2, db_builder.cpp:392-458) against one global sort, and the sharded
    writer against the single-database writer byte for byte"""
    from ipk_b200 import hostlib
    rng = np.random.default_rng(40 + n_shards)
    shards = _random_shards(rng, n_shards, 70000)
    keys, fv, offsets, branch, score = hostlib.merge_shards(shards, threads)
    all_keys = np.concatenate([s.keys for s in shards]); all_fv = np.concatenate([s.fv for s in shards])
    img = all_fv.view(np.uint32).astype(np.int64)
    img = np.where(img >> 31, img ^ 0xffffffff, img ^ 0x80000000)            # the device's order of floats: -0.0 before 0.0
    o = np.lexsort((all_keys, img))
    assert np.array_equal(keys, all_keys[o]) and np.array_equal(fv.view(np.uint32), all_fv[o].view(np.uint32))
    # entries follow their k-mer
    k0 = int(keys[12345])
    s0 = next(s for s in shards if k0 in set(s.keys[:].tolist()))
    i0 = int(np.where(s0.keys == k0)[0][0])
    a, b = int(offsets[12345]), int(offsets[12346])
    assert np.array_equal(branch[a:b], s0.branch[int(s0.offsets[i0]):int(s0.offsets[i0 + 1])])
    one, many = str(tmp_path / "one.ipk"), str(tmp_path / "many.ipk")
    hostlib.write_ipk(one, keys, fv, offsets, branch, score, 9, 1.5, "DNA", "(a:1,b:2);", [1, 1, 3], [0, 0, 3])
    hostlib.write_ipk_sharded(many, shards, 9, 1.5, "DNA", "(a:1,b:2);", [1, 1, 3], [0, 0, 3], threads)
    assert open(one, "rb").read() == open(many, "rb").read()


def test_exchange_plan_arithmetic():
    """ipkgpu_exchange_plan (host-only, no device): owners are contiguous balanced bucket ranges, the bases tile every owner's
    buffer exactly -- bucket after bucket, source rank after source rank"""
    from ipk_b200 import capi
    rng = np.random.default_rng(8)
    for world in (1, 2, 3, 8):
        counts = rng.integers(0, 5000, (world, 256)).astype(np.uint64)
        counts[:, rng.integers(0, 256, 20)] = 0
        plans = [capi.exchange_plan(world, r, counts) for r in range(world)]
        owner = plans[0][0]
        total = counts.sum(0)
        assert all(np.array_equal(owner, p[0]) for p in plans) and np.all(np.diff(owner.astype(int)) >= 0) and owner.max() < world
        load = np.array([int(total[owner == r].sum()) for r in range(world)])
        assert [p[3] for p in plans] == list(load)
        assert load.max() - load.min() <= 2 * int(total.max())                 # balanced to a bucket
        for o in range(world):
            spans = sorted((int(plans[r][1][d]), int(plans[r][1][d]) + int(counts[r, d])) for r in range(world) for d in range(256) if owner[d] == o and counts[r, d])
            assert spans[0][0] == 0 and spans[-1][1] == load[o] and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            start = plans[o][2]
            assert start[256] == load[o] and all(int(start[d + 1]) - int(start[d]) == (int(total[d]) if owner[d] == o else 0) for d in range(256))


@pytest.mark.gpu
def test_ipk_aa_pos_executable(tmp_path):
    """`ipk-aa-pos` (the KEEP_POSITIONS binary of the reference, picked by name): AR file + node map -> DB.ipk whose entries carry
    the position of the first window that reached their score (branch_group.cpp:72-86)"""
    import oraclelib as O
    from test_gpu_parity import _positions_oracle
    from ipk_b200 import hostlib, capi, synth
    n_taxa, W, sigma, k, omega = 4, 50, 20, 3, 1.5
    logp, nb = synth.posteriors(n_taxa, W, sigma, seed=5)
    order = "RHKDESTNQCGPAILMFWYV"
    probs = np.power(10.0, logp.astype(np.float64))[:, :, [order.index(c) for c in AA_RAXML]]
    labels = [f"Node{i + 1}" for i in range(len(nb))]
    tsv, nmap, out = str(tmp_path / "ar.raxml.ancestralProbs"), str(tmp_path / "node_map.tsv"), str(tmp_path / "DB.ipk")
    write_tsv(tsv, probs, labels, sigma)
    with open(nmap, "w") as f:
        f.write("".join(f"{lab}\t{int(b)}\n" for lab, b in zip(labels, nb)))
    exe = os.path.join(os.path.dirname(hostlib.EXE_PATH), "ipk-aa-pos")
    r = subprocess.run([exe, "--ar-probs", tsv, "--node-map", nmap, "-k", str(k), "--omega", str(omega), "--mif0", "-o", out,
                        "--node-count", str(2 * n_taxa - 1), "--merge-branches"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    d = hostlib.read_ipk(out)
    assert d["sequence_type"] == "Proteins" and d["position"] is not None
    _, parsed = hostlib.read_ar(tsv, sigma, labels)
    thr, lthr = capi.score_threshold(omega, sigma, k)
    want = _positions_oracle(parsed, nb, sigma, 5, k, lthr)
    keys = np.repeat(d["keys"], np.diff(d["offsets"].astype(np.int64)))
    have = {(int(a), int(b)): (int(s), int(p)) for a, b, s, p in zip(keys, d["branch"], d["score"].view(np.uint32), d["position"])}
    assert have == want and len(want) > 100
