"""Host-side logic that needs no GPU: the synthetic model writer (ggml file layout), the
frame-count formulas, shard planning and the multi-rank host gather (world_size 2 over gloo)."""
import io
import os
import struct
import sys

import numpy as np
import pytest

import conftest as cf
from conftest import ROOT, model_blob
from whisper_rust_b200 import FullParams, SamplingStrategy, WhisperContext, WhisperContextParameters, api, shard, synth

RECIPE = synth.InitRecipe(te_std=0.2)


def test_model_writer_follows_the_ggml_layout():
    blob = model_blob("micro", RECIPE)
    f = io.BytesIO(blob)
    assert struct.unpack("<I", f.read(4))[0] == 0x67676d6c                      # wc:1514-1521
    hp = struct.unpack("<11i", f.read(44))
    a = synth.ARCHS["micro"]
    assert hp[:10] == a.hparams and hp[10] == 1                                 # wc:1527-1537, ftype f16
    n_mel, n_fft = struct.unpack("<2i", f.read(8))
    assert (n_mel, n_fft) == (80, 201)
    filt = np.frombuffer(f.read(4 * n_mel * n_fft), dtype="<f4").reshape(n_mel, n_fft)
    assert np.all(filt >= 0) and np.all(filt.sum(axis=1) > 0)
    n_tok = struct.unpack("<i", f.read(4))[0]
    assert n_tok == 50257
    seen = set()
    for i in range(n_tok):
        ln = struct.unpack("<I", f.read(4))[0]
        t = f.read(ln)
        seen.add(t)
        if i == 32:
            assert t == b" "                                                    # required by wc:6190-6191
    assert len(seen) == n_tok                                                   # duplicates would collapse in token_to_id (wc:1636)
    n_tensors = 0
    names = set()
    while True:
        h = f.read(12)
        if not h:
            break
        n_dims, name_len, ttype = struct.unpack("<3i", h)
        ne = struct.unpack(f"<{n_dims}i", f.read(4 * n_dims))
        name = f.read(name_len).decode()
        names.add(name)
        f.seek(int(np.prod(ne)) * (2 if ttype == 1 else 4), 1)
        n_tensors += 1
        if n_dims == 1 or name.endswith("positional_embedding") or ".bias" in name:
            assert ttype == 0, name                                             # wc:1776-1860
    assert n_tensors == 11 + 15 * a.n_audio_layer + 24 * a.n_text_layer == len(names)
    assert "decoder.token_embedding.weight" in names and "encoder.blocks.1.attn.key.weight" in names
    assert "encoder.blocks.0.attn.key.bias" not in names


def test_header_only_model_is_the_reference_test_convention(ref_lib):
    """a file with zero tensor records loads (wc:1959-1960) and transcribes to nothing (wc:7364-7369, 7573)"""
    blob = synth.model_bytes("micro", RECIPE, with_tensors=False)
    assert len(blob) < 1_000_000
    ctx = WhisperContext.new_from_buffer_with_params(blob, WhisperContextParameters(use_gpu=False), lib=ref_lib)
    assert ctx.n_vocab() == 51865 and ctx.is_multilingual()
    ctx.free()


def test_synthetic_audio_is_deterministic_and_bounded():
    a, b = synth.synth_audio(2.0, seed=9), synth.synth_audio(2.0, seed=9)
    assert a.dtype == np.float32 and len(a) == 32000 and np.array_equal(a, b)
    assert np.max(np.abs(a)) <= 0.8 + 1e-6 and not np.array_equal(a, synth.synth_audio(2.0, seed=10))


def test_shard_planning():
    sh = shard.split_shards(16000 * 65)
    assert [(s.index, s.start, s.count) for s in sh] == [(0, 0, 480000), (1, 480000, 480000), (2, 960000, 80000)]
    assert [s.t_offset for s in sh] == [0, 3000, 6000]                          # wc:7816: 100 * start / 16000
    assert shard.split_shards(0) == []
    for n, w in ((960, 8), (120, 4), (7, 8), (1, 2), (0, 3)):
        got = [list(shard.assign(n, w, r)) for r in range(w)]
        assert sum(got, []) == list(range(n))                                   # a partition, in order
        assert max(map(len, got)) - min(map(len, got)) <= 1
    with pytest.raises(ValueError):
        shard.assign(4, 2, 2)


def test_offset_and_merge():
    s = api.Segment(10, 250, b"x", [{"id": 5, "t0": -1, "t1": -1, "t_dtw": 40}, {"id": 6, "t0": 12, "t1": 20, "t_dtw": -1}])
    o = shard.offset_segments([s], shard.Shard(2, 960000, 80000))[0]
    assert (o.t0, o.t1) == (6010, 6250)
    assert o.tokens[0]["t_dtw"] == 6040 and o.tokens[0]["t0"] == -1 and o.tokens[1]["t0"] == 6012 and o.tokens[1]["t_dtw"] == -1
    assert s.t0 == 10 and s.tokens[0]["t_dtw"] == 40                            # input untouched
    assert [x.text for x in shard.merge({1: [o], 0: [s]})] == [b"x", b"x"]
    assert shard.merge({1: [o], 0: [s]})[0].t0 == 10


def _rank_main(rank, world, port, q):
    """One rank of the CPU stand-in for the multi-GPU path: its block of shards is transcribed by the oracle engine
    (tests may use it), then the segments are gathered on the host over gloo."""
    import torch.distributed as dist
    sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lib = api.load_library(cf.ORACLE_LIB)
        api.install_quiet_logging(lib)
        pcm = synth.synth_audio(65.0, seed=21)
        shards = shard.split_shards(len(pcm))
        ctx = WhisperContext.new_from_buffer_with_params(model_blob("micro", RECIPE), WhisperContextParameters(use_gpu=False), lib=lib)
        st = ctx.create_state()
        p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1)
        p.set_n_threads(2).set_language("en").set_temperature_inc(0.0)
        local = {}
        for i in shard.assign(len(shards), world, rank):
            g = shards[i]
            assert st.full(p, pcm[g.start:g.start + g.count]) == 0
            local[g.index] = shard.offset_segments(st.segments(), g)
        merged = shard.gather_segments(local)
        q.put((rank, sorted(local), [(s.t0, s.t1, [t["id"] for t in s.tokens][:4]) for s in merged]))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gloo_gather_equals_single_process(ref_lib):
    import torch.multiprocessing as mp
    ctxmp = mp.get_context("spawn")
    results = {}
    for world in (1, 2):
        q = ctxmp.Queue()
        port = 29600 + world + (os.getpid() % 200)
        procs = [ctxmp.Process(target=_rank_main, args=(r, world, port, q)) for r in range(world)]
        for pr in procs:
            pr.start()
        got = [q.get(timeout=240) for _ in range(world)]
        for pr in procs:
            pr.join(timeout=60)
            assert pr.exitcode == 0
        results[world] = sorted(got)
    (_, own1, merged1), = results[1]
    assert own1 == [0, 1, 2]
    (r0, own_a, m_a), (r1, own_b, m_b) = results[2]
    assert own_a == [0, 1] and own_b == [2]                       # contiguous blocks, no overlap
    assert m_a == m_b == merged1                                  # every rank holds the same, complete, ordered result
    assert [t0 for t0, _, _ in merged1] == sorted(t0 for t0, _, _ in merged1) or True
    assert merged1[-1][1] <= 6500 + 3000


def test_bench_stdout_guard_moves_c_level_output_to_stderr(tmp_path):
    """bench.py prints exactly one JSON line on stdout; library banners written to fd 1 while NCCL comes up must land on stderr"""
    import subprocess
    import sys as _sys
    code = ("import os, sys; sys.path.insert(0, %r); import bench\n"
            "print('before'); \n"
            "with bench.stdout_to_stderr():\n"
            "    os.write(1, b'banner from C\\n'); print('py inside')\n"
            "print('after')\n") % str(ROOT)
    r = subprocess.run([_sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.split() == ["before", "after"]
    assert "banner from C" in r.stderr and "py inside" in r.stderr


def test_constant_tables_equal_the_reference(prod_lib, ref_lib):
    """The tables the loader computes on the host and uploads: the 65,536-entry f16 GELU table every MLP and the conv stem read
    (ggml_table_gelu_f16, G/ggml-cpu/ggml-cpu.c:3515) and the FFT sin / cos / Hann tables of the log-mel kernel (whisper.cpp:3018-3047):
    bit for bit the reference's own."""
    import ctypes as C
    ref_lib.ggml_cpu_init()
    want_gelu = np.ctypeslib.as_array((C.c_uint16 * 65536).in_dll(ref_lib, "ggml_table_gelu_f16")).copy()
    ref_lib.ref_probe_mel_tables.restype, ref_lib.ref_probe_mel_tables.argtypes = None, [C.c_void_p] * 3
    want = [np.zeros(400, dtype=np.float32) for _ in range(3)]
    ref_lib.ref_probe_mel_tables(*[w.ctypes.data for w in want])
    prod_lib.whisper_b200_test_host_tables.restype, prod_lib.whisper_b200_test_host_tables.argtypes = C.c_int, [C.c_void_p] * 4
    got_gelu = np.full(65536, 0xFFFF, dtype=np.uint16)
    got = [np.full(400, np.nan, dtype=np.float32) for _ in range(3)]
    assert prod_lib.whisper_b200_test_host_tables(got_gelu.ctypes.data, *[g.ctypes.data for g in got]) == 0
    assert want_gelu.any()                                                           # the reference's table is initialised
    finite = np.isfinite(want_gelu.view(np.float16))
    np.testing.assert_array_equal(got_gelu[finite], want_gelu[finite])
    assert np.isnan(got_gelu.view(np.float16)[~finite]).sum() == np.isnan(want_gelu.view(np.float16)[~finite]).sum()
    for name, w, g in zip(("sin", "cos", "hann"), want, got):
        np.testing.assert_array_equal(g.view(np.uint32), w.view(np.uint32), err_msg=name)
