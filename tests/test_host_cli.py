"""The C++14 host program `charon_b200/charon` (dehost / classify CLI over libcharon_gpu.so).

CPU part: command line handling and the `.idx` loader (no GPU needed).
GPU part: whole-process parity — an index file + FASTQ/FASTA on disk, `charon dehost|classify` run as a subprocess, its
stdout compared byte for byte with the oracle's restatement of the reference's Result/print path
(include/result.hpp:97-213, include/read_entry.hpp:322-337) at -t 1 semantics.
"""
import gzip
import os
import subprocess

import numpy as np
import pytest

import oracle
from tests import common

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHARON = os.path.join(ROOT, "charon_b200", "charon")


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "charon_b200", "csrc")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "charon_b200", "host")])


@pytest.fixture(scope="module")
def charon():
    if not os.path.exists(CHARON):
        _build()
    return CHARON


def run(charon, *args, cwd=None):
    return subprocess.run([charon, *args], cwd=cwd, capture_output=True, text=True, timeout=600)


# ------------------------------------------------------------------------------------------- CPU
def test_version_and_help(charon):
    r = run(charon, "--version")
    assert r.returncode == 0 and "charon" in r.stdout
    r = run(charon, "dehost", "--help")
    assert r.returncode == 0
    for opt in ("--db", "-e,--extract", "-p,--prefix", "--chunk_size", "--lo_hi_threshold", "--num_reads_to_fit", "-d,--dist", "--min_length",
                "--min_quality", "--min_compression", "--confidence", "--host_unique_prop_lo_threshold", "--min_proportion_diff",
                "--min_probability_diff", "--log", "-t,--threads", "-v"):
        assert opt in r.stdout, opt
    r = run(charon, "classify", "--help")
    assert "--host_unique_prop_lo_threshold" not in r.stdout and "--min_probability_diff" not in r.stdout


def test_fastq_reader_fast_paths(tmp_path):
    """CPU: records mixing plain runs with lower case, IUPAC codes, U and N come out as the dna5 upper-case string whatever the
    alignment of the 16-letter fast path; a quality character outside '!' .. '~' is an error."""
    exe = str(tmp_path / "fastx_check")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O2", "-std=c++14", os.path.join(ROOT, "tests", "fastx_check.cpp"), "-o", exe, "-lz"])
    out = subprocess.run([exe, str(tmp_path / "r.fq")], capture_output=True, text=True)
    assert out.returncode == 0 and "bad 0" in out.stdout, out.stdout + out.stderr


def test_cli_errors(charon, tmp_path):
    fq = tmp_path / "r.fq"
    fq.write_text("@r1\nACGT\n+\nIIII\n")
    assert run(charon, "dehost").returncode != 0                                         # <fastaq> required
    assert run(charon, "dehost", str(fq)).returncode != 0                                # --db required
    assert run(charon, "dehost", str(fq), "--db", str(tmp_path / "nope")).returncode != 0  # ExistingPath
    assert run(charon, "dehost", str(tmp_path / "missing.fq"), "--db", str(fq)).returncode != 0
    assert run(charon, "dehost", str(fq), "--db", str(fq), "--confidence", "300").returncode != 0   # uint8_t option
    assert run(charon, "dehost", str(fq), "--db", str(fq), "--bogus").returncode != 0
    assert run(charon, "frobnicate").returncode != 0


def test_inspect_reads_the_oracle_written_index(charon, tmp_path):
    idx, _ = common.build_oracle_index([60_000, 40_000, 30_000], [0, 1, 1], ["host", "microbial"], seed=11)
    p = str(tmp_path / "t.idx")
    idx.save(p)
    r = run(charon, "inspect", p)
    assert r.returncode == 0, r.stderr
    kv = {l.split("\t")[0]: l.split("\t")[1:] for l in r.stdout.strip().split("\n")}
    assert kv["window_size"] == ["41"] and kv["kmer_size"] == ["19"] and kv["num_bins"] == ["3"]
    assert kv["categories"] == ["host", "microbial"]
    assert kv["bin_to_cat"] == ["0", "1", "1"]
    assert f"bin_size={idx.bin_size}" in kv["ibf"][0] and "hash_funs=3" in kv["ibf"][0] and "bins=3 " in kv["ibf"][0]
    ones = int(np.unpackbits(idx.rows().view(np.uint8)).sum())
    assert f"ones={ones}" in kv["sd_vector"][0]
    assert kv["host_category_index"] == ["0"]


def test_inspect_rejects_corrupt_index(charon, tmp_path):
    idx, _ = common.build_oracle_index([30_000, 30_000], [0, 1], ["host", "microbial"], seed=12)
    p = str(tmp_path / "t.idx")
    idx.save(p)
    raw = bytearray(open(p, "rb").read())
    trunc = str(tmp_path / "trunc.idx")
    open(trunc, "wb").write(raw[:200])
    assert run(charon, "inspect", trunc).returncode != 0
    bad = bytearray(raw)
    # IBF header follows the summaries; flip hash_shift (search for the 6-u64 header by its bins/technical_bins pair)
    key = np.array([2, 64, idx.bin_size], dtype=np.uint64).tobytes()
    off = bytes(bad).find(key)
    assert off > 0
    bad[off + 24] ^= 1
    badp = str(tmp_path / "bad.idx")
    open(badp, "wb").write(bad)
    r = run(charon, "inspect", badp)
    assert r.returncode != 0 and "hash_shift" in r.stderr


# ------------------------------------------------------------------------------------------- GPU
def _mixed_reads(refs, n, seed, min_len=10, max_len=6000):
    """Reads of mixed length from the references (both strands untouched; content only matters for parity), with a
    few lower-case / IUPAC / N characters, plus qualities."""
    rng = np.random.default_rng(seed)
    reads = []
    for i in range(n):
        L = int(rng.integers(min_len, max_len))
        src = int(rng.integers(0, len(refs) + 1))
        if src < len(refs):
            s = int(rng.integers(0, len(refs[src]) - L))
            seq = refs[src][s:s + L].copy()
            mut = rng.random(L) < 0.02
            seq[mut] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(mut.sum()))]
        else:
            seq = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, L)].copy()
        if i % 7 == 0:
            seq[:: max(3, L // 5)] = ord("N")
        text = seq.tobytes().decode()
        if i % 11 == 0:
            text = text.lower()
        if i % 13 == 0:
            text = text[: L // 2] + "RYKM" + text[L // 2 + 4:] if L > 8 else text
        qual = "".join(chr(33 + int(q)) for q in rng.integers(5, 40, len(text)))
        reads.append((f"read{i} extra words {i}", text, qual))
    return reads


def _norm(text):   # dna5 upper-case text as the reference's sequence_to_string gives (src/utils.cpp:105-112)
    t = text.upper().replace("U", "T")
    return "".join(c if c in "ACGT" else "N" for c in t)


def _scalars(reads_groups):
    mq, cr = [], []
    for group in reads_groups:
        quals = "".join(q for _, _, q in group)
        s = sum(ord(c) - 33 for c in quals)
        mq.append(np.float32(s) / np.float32(len(quals)) if quals else np.float32(0))
        cr.append(oracle.compression_ratio("".join(_norm(t) for _, t, _ in group)))
    return np.array(mq, np.float32), np.array(cr, np.float32)


def _write_fastq(path, reads, gz=False):
    data = "".join(f"@{i}\n{s}\n+\n{q}\n" for i, s, q in reads)
    if gz:
        with gzip.open(path, "wt") as f:
            f.write(data)
    else:
        open(path, "w").write(data)


def _oracle_tsv(idx, groups, opts):
    """groups: list of 1- or 2-tuples of (id, seq, qual)."""
    pieces, seg_off, read_seg = [], [0], [0]
    for g in groups:
        for _, t, _ in g:
            pieces.append(np.frombuffer(t.encode(), np.uint8))
            seg_off.append(seg_off[-1] + len(t))
        read_seg.append(len(seg_off) - 1)
    bases = np.concatenate(pieces)
    mq, cr = _scalars(groups)
    ids = [g[0][0] for g in groups]
    return idx.run_tsv(ids, bases, np.array(seg_off, np.uint64), mq, cr, read_seg=np.array(read_seg, np.uint32), opts=opts)


@pytest.fixture(scope="module")
def small_index(tmp_path_factory):
    d = tmp_path_factory.mktemp("idx")
    idx, refs = common.build_oracle_index([300_000, 250_000], [0, 1], ["host", "microbial"], seed=21)
    p = str(d / "small.idx")
    idx.save(p)
    return idx, refs, p


@pytest.mark.gpu
@pytest.mark.parametrize("gz,threads", [(False, 1), (True, 4)])
def test_dehost_stdout_is_byte_identical(charon, small_index, tmp_path, gz, threads):
    idx, refs, idx_path = small_index
    reads = _mixed_reads(refs, 400, seed=5)
    reads.insert(3, ("tiny", "ACGTACG", "IIIIIII"))          # shorter than k: num_hashes 0 -> nan columns
    reads.insert(9, ("empty", "", ""))                        # zero length: skipped (src/dehost_main.cpp:351-354)
    fq = str(tmp_path / ("r.fq.gz" if gz else "r.fq"))
    _write_fastq(fq, reads, gz)
    want = _oracle_tsv(idx, [(r,) for r in reads], oracle.default_options(True))
    got = run(charon, "dehost", fq, "--db", idx_path, "-t", str(threads), "--log", str(tmp_path / "c.log"), "--batch_reads", "97", cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == want
    assert want.count("\n") == len(reads) - 2                 # first read dropped (Appendix B-1) + the empty one
    assert "nan" in want
    assert os.path.exists(tmp_path / "c.log")


@pytest.mark.gpu
def test_dehost_fasta_input_and_db_prefix(charon, small_index, tmp_path):
    idx, refs, idx_path = small_index
    reads = [(i, s, "") for i, s, _ in _mixed_reads(refs, 60, seed=6, min_len=200)]
    fa = str(tmp_path / "r.fa")
    with open(fa, "w") as f:
        for i, s, _ in reads:
            f.write(f">{i}\n")
            for o in range(0, len(s), 60):
                f.write(s[o:o + 60] + "\n")
    want = _oracle_tsv(idx, [(r,) for r in reads], oracle.default_options(True))
    got = run(charon, "dehost", fa, "--db", idx_path, cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == want            # mean quality 0 < min_quality: every row U, still printed


@pytest.mark.gpu
def test_dehost_paired_uses_call_category(charon, small_index, tmp_path):
    idx, refs, idx_path = small_index
    a = _mixed_reads(refs, 120, seed=7, min_len=60, max_len=300)
    b = _mixed_reads(refs, 120, seed=8, min_len=60, max_len=300)
    r1 = [(f"pair{i}/1", s, q) for i, (_, s, q) in enumerate(a)]
    r2 = [(f"pair{i}/2", s, q) for i, (_, s, q) in enumerate(b)]
    f1, f2 = str(tmp_path / "r_1.fq"), str(tmp_path / "r_2.fq")
    _write_fastq(f1, r1); _write_fastq(f2, r2)
    o = oracle.default_options(True)
    o.dehost = 0          # add_paired_read(read, r1, r2) -> dehost=false -> call_category (src/dehost_main.cpp:470)
    o.min_length = 80     # src/dehost_main.cpp:493-496
    want = _oracle_tsv(idx, list(zip(r1, r2)), o)
    got = run(charon, "dehost", f1, f2, "--db", idx_path, cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == want
    # mismatching pair ids stop the run with the reference's message
    _write_fastq(f2, [("other/2",) + r2[0][1:]] + r2[1:])
    bad = run(charon, "dehost", f1, f2, "--db", idx_path, cwd=tmp_path)
    assert bad.returncode != 0 and "pairs don't match" in bad.stderr


@pytest.mark.gpu
def test_dehost_extract_trains_then_classifies(charon, small_index, tmp_path):
    idx, refs, idx_path = small_index
    reads = _mixed_reads(refs, 600, seed=9, min_len=400, max_len=3000)
    fq = str(tmp_path / "r.fastq")
    _write_fastq(fq, reads)
    o = oracle.default_options(True)
    o.run_extract = 1
    o.num_reads_to_fit = 25      # cache capacity 25 * 2 * 4 = 200 reads (include/result.hpp:84)
    want = _oracle_tsv(idx, [(r,) for r in reads], o)
    got = run(charon, "dehost", fq, "--db", idx_path, "-e", "microbial", "-p", "out", "--num_reads_to_fit", "25", "--batch_reads", "64", cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == want
    called = [l.split("\t")[1] for l in want.strip().split("\n") if l.split("\t")[2] == "microbial"]
    assert called
    with gzip.open(tmp_path / "out_microbial.fastq.gz", "rt") as f:
        lines = f.read().split("\n")
    assert [l[1:].split(" ")[0] for l in lines[0::4] if l] == called
    by_id = {i.split(" ")[0]: (s, q) for i, s, q in reads}
    assert lines[1] == _norm(by_id[called[0]][0]) and lines[3] == by_id[called[0]][1]


@pytest.mark.gpu
def test_gpu_compression_flag_gives_the_same_bytes(charon, small_index, tmp_path):
    """--gpu_compression: the compression ratio column and every call that depends on it come from gzip_size_kernel (zlib
    level-6 gzip size, bit-exact) instead of zlib on the host threads: stdout must not change — single reads (with a tiny, an
    empty, a low-complexity and a 70 000-letter read, beyond zlib's 64 KB window), paired reads, and --extract."""
    idx, refs, idx_path = small_index
    reads = _mixed_reads(refs, 300, seed=15)
    reads.insert(3, ("tiny", "ACGTACG", "IIIIIII"))
    reads.insert(9, ("empty", "", ""))
    reads.insert(20, ("polyA", "A" * 2000, "I" * 2000))
    rng = np.random.default_rng(3)
    big = "".join(rng.choice(list("ACGT"), 70_000))
    reads.insert(150, ("longer_than_the_window", big, "I" * len(big)))
    fq = str(tmp_path / "r.fq")
    _write_fastq(fq, reads)
    want = _oracle_tsv(idx, [(r,) for r in reads], oracle.default_options(True))
    for flag in ([], ["--gpu_compression"]):
        got = run(charon, "dehost", fq, "--db", idx_path, "-t", "2", "--batch_reads", "50", *flag, cwd=tmp_path)
        assert got.returncode == 0, got.stderr
        assert got.stdout == want, flag
    # paired
    a = _mixed_reads(refs, 90, seed=17, min_len=60, max_len=300)
    b = _mixed_reads(refs, 90, seed=18, min_len=60, max_len=300)
    r1 = [(f"pair{i}/1", s, q) for i, (_, s, q) in enumerate(a)]
    r2 = [(f"pair{i}/2", s, q) for i, (_, s, q) in enumerate(b)]
    f1, f2 = str(tmp_path / "p_1.fq"), str(tmp_path / "p_2.fq")
    _write_fastq(f1, r1); _write_fastq(f2, r2)
    o = oracle.default_options(True)
    o.dehost = 0
    o.min_length = 80
    got = run(charon, "dehost", f1, f2, "--db", idx_path, "--gpu_compression", "--batch_reads", "40", cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == _oracle_tsv(idx, list(zip(r1, r2)), o)
    # --extract: training batches keep the host's ratios, the rest use the kernel's
    reads = _mixed_reads(refs, 500, seed=19, min_len=400, max_len=3000)
    fq = str(tmp_path / "e.fastq")
    _write_fastq(fq, reads)
    o = oracle.default_options(True)
    o.run_extract = 1
    o.num_reads_to_fit = 25
    got = run(charon, "dehost", fq, "--db", idx_path, "-e", "microbial", "-p", "outg", "--num_reads_to_fit", "25", "--batch_reads", "64", "--gpu_compression", cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == _oracle_tsv(idx, [(r,) for r in reads], o)


@pytest.mark.gpu
def test_classify_three_categories(charon, tmp_path):
    idx, refs = common.build_oracle_index([200_000, 150_000, 120_000, 90_000], [0, 1, 1, 2], ["bacteria", "host", "virus"], seed=31)
    p = str(tmp_path / "three.idx")
    idx.save(p)
    reads = _mixed_reads(refs, 300, seed=10, min_len=150, max_len=4000)
    fq = str(tmp_path / "r.fq")
    _write_fastq(fq, reads)
    want = _oracle_tsv(idx, [(r,) for r in reads], oracle.default_options(False))
    got = run(charon, "classify", fq, "--db", p, cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    wl, gl = want.strip().split("\n"), got.stdout.strip().split("\n")
    assert len(wl) == len(gl)
    for w, g in zip(wl, gl):
        wf, gf = w.split("\t"), g.split("\t")
        assert wf[1] == gf[1] and wf[3:8] == gf[3:8]          # id, length, num_hashes, quality, confidence, compression
        for wc, gc in zip(wf[8].split(), gf[8].split()):
            wp, gp = wc.split(":"), gc.split(":")
            assert wp[:4] == gp[:4]                           # name, count, proportion, unique proportion: exact
            # beta densities go through gcem in the reference (absent): best effort, SURVEY A.5
            assert wp[4] == gp[4] or abs(float(wp[4]) - float(gp[4])) <= 1e-4 * max(abs(float(wp[4])), 1e-30)
    same_call = sum(w.split("\t")[:3] == g.split("\t")[:3] for w, g in zip(wl, gl))
    assert same_call >= len(wl) - 2


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [[], ["--partition_filter"]])
def test_dehost_two_gpus_same_bytes(charon, small_index, tmp_path, extra):
    if _n_gpus() < 2:
        pytest.skip("needs two GPUs")
    idx, refs, idx_path = small_index
    reads = _mixed_reads(refs, 500, seed=15)
    fq = str(tmp_path / "r.fq")
    _write_fastq(fq, reads)
    want = _oracle_tsv(idx, [(r,) for r in reads], oracle.default_options(True))
    got = run(charon, "dehost", fq, "--db", idx_path, "--gpus", "2", "--batch_reads", "60", *extra, cwd=tmp_path)
    assert got.returncode == 0, got.stderr
    assert got.stdout == want


def _write_fasta(path, records, width=70):
    with open(path, "w") as f:
        for name, seq in records:
            f.write(f">{name}\n")
            for o in range(0, len(seq), width):
                f.write(seq[o:o + width] + "\n")


@pytest.mark.gpu
def test_charon_index_writes_the_same_filter_as_the_oracle(charon, tmp_path):
    """`charon index` (GPU hashing, distinct counts, insertion, Elias-Fano encoding) -> .idx; reloaded by the oracle's reader it
    must hold exactly the bits and statistics of the oracle's restatement of src/index_main.cpp on the same files."""
    from tools import synth
    lens = [(90_000, 30_000), (120_000,), (50_000, 20_000, 10_000)]
    cats = ["host", "microbial", "microbial"]
    seqs, lines = [], []
    for b, parts in enumerate(lens):
        recs = [(f"chr{b}_{j}", synth.reference(300 + 10 * b + j, n).tobytes().decode()) for j, n in enumerate(parts)]
        p = str(tmp_path / f"ref{b}.fa")
        _write_fasta(p, recs)
        seqs.append([s for _, s in recs])
        lines.append(f"{p}\t{cats[b]}")
    tab = tmp_path / "input.tsv"
    tab.write_text("\n".join(lines) + "\n")
    r = run(charon, "index", str(tab), "-p", str(tmp_path / "mine"), cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    got = oracle.Index.load(str(tmp_path / "mine.idx"))
    records = [np.frombuffer(s.encode(), np.uint8) for rs in seqs for s in rs]
    rec_bin = [b for b, rs in enumerate(seqs) for _ in rs]
    want = oracle.Index.build(records, rec_bin=rec_bin, bin_cat=[0, 1, 1], categories=["host", "microbial"])
    assert (got.k, got.w, got.bins, got.bin_size, got.hash_funs) == (want.k, want.w, want.bins, want.bin_size, want.hash_funs)
    assert np.array_equal(got.rows(), want.rows())
    assert list(got.bin_to_cat) == [0, 1, 1]
    insp = run(charon, "inspect", str(tmp_path / "mine.idx"))
    assert insp.returncode == 0, insp.stderr
    # and the product reads its own file: dehost against the oracle on the oracle's own index
    reads = _mixed_reads([np.frombuffer(s.encode(), np.uint8) for s in (seqs[0][0], seqs[1][0])], 120, seed=33, min_len=300, max_len=2500)
    fq = str(tmp_path / "r.fq")
    _write_fastq(fq, reads)
    out = run(charon, "dehost", fq, "--db", str(tmp_path / "mine.idx"), cwd=tmp_path)
    assert out.returncode == 0, out.stderr
    assert out.stdout == _oracle_tsv(want, [(x,) for x in reads], oracle.default_options(True))


@pytest.mark.gpu
def test_charon_index_optimize_merges_bins(charon, tmp_path):
    from tools import synth
    sizes = [100_000, 20_000, 15_000, 10_000, 60_000]
    cats = ["host", "microbial", "microbial", "microbial", "host"]
    lines = []
    for b, n in enumerate(sizes):
        p = str(tmp_path / f"g{b}.fasta")
        _write_fasta(p, [(f"g{b}", synth.reference(500 + b, n).tobytes().decode())])
        lines.append(f"{p}\t{cats[b]}")
    tab = tmp_path / "in.tsv"
    tab.write_text("\n".join(lines) + "\n")
    assert run(charon, "index", str(tab), cwd=tmp_path).returncode == 0          # default output <input>.idx
    plain = oracle.Index.load(str(tab) + ".idx")
    r = run(charon, "index", str(tab), "-p", str(tmp_path / "opt"), "--optimize", cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    opt = oracle.Index.load(str(tmp_path / "opt.idx"))
    assert plain.bins == 5 and opt.bins < plain.bins                              # the three small microbial files share buckets
    assert sorted(set(opt.bin_to_cat)) == [0, 1]
    # merged filter answers "present" wherever any merged bin did: per category the OR of the plain masks
    vals = oracle.minimisers(synth.reference(501, 20_000), 19, 41)[:2000]
    pm, om = plain.query(vals)[:, 0], opt.query(vals)[:, 0]
    for c in (0, 1):
        pbits = sum(1 << b for b in range(plain.bins) if plain.bin_to_cat[b] == c)
        obits = sum(1 << b for b in range(opt.bins) if opt.bin_to_cat[b] == c)
        assert np.all(((pm & np.uint64(pbits)) != 0) <= ((om & np.uint64(obits)) != 0))
