"""mio_peak_normalize_device_noalloc, the one entry point of the library that allocates nothing.

Some users of this library are not allowed to allocate device memory while they run: a plug-in inside a serving process
whose own pool claimed all of HBM at start-up, or a real-time audio thread that must never queue behind the driver's
allocator lock.  For them the per-utterance peak normalisation exists in a form whose scratch (one float per utterance) is a
fixed array inside the library image instead of a context buffer.  This file checks that form against the ordinary path
(peak_normalize = 1 inside mio_decode_wave_batch), through the C ABI, bit for bit -- the same bar the batch-invariance tests
hold the ordinary path to.

The array is shared by all calls and keeps the previous call's maxima, so the test makes several calls in a row and
arranges for a stale maximum to be visible: the first call sees the waveform scaled by 4 (exact in floating point, and the
normalised result is unchanged), the second the unscaled one -- if the slots were not cleared the second result would be 4x
too quiet.

Named test_zz_* so that pytest collects it after every other file: `pytest -x` stops at the first failure, and a problem
here must not hide the parity tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(seed, lens, voice_ids):
    from miocodec_b200 import synth
    cu = np.cumsum([0] + lens).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), seed)
    vt = np.stack([synth.synth_voice(1), synth.synth_voice(2)]).astype(np.float32)
    return codes, cu, vt, np.array(voice_ids, dtype=np.int32)


def _resident_unnormalised(codec, batch):
    """Decode without normalisation and leave the waveform in HBM: (torch view of it, int64 sample offsets on the device)."""
    import torch
    from miocodec_b200 import sharding
    codes, cu, vt, vids = batch
    codec.batch_upload(codes, cu, vt, vids)
    codec.batch_run(False)
    codec.batch_sync()
    ptr, n = codec.wave_device()
    wave = sharding.device_tensor(ptr, n, "float32", 0)
    offs = torch.tensor(cu.astype(np.int64) * codec.samples_per_token, dtype=torch.int64, device="cuda:0")
    return wave, offs


def test_noalloc_peak_normalize_matches_the_context_path(codec):
    import torch
    from miocodec_b200 import binding
    stream = torch.cuda.current_stream().cuda_stream

    a = _batch(311, [33, 17, 60], [0, 1, 0])
    want_a = codec.decode_wave_batch(a[0], a[1], a[2], a[3], peak_normalize=True)      # the existing entry point
    assert 0.94 < np.abs(want_a).max() <= 0.9501
    wave, offs = _resident_unnormalised(codec, a)
    n_utts = len(a[1]) - 1

    # call 1, out of place, on the waveform made 4x louder: leaves maxima 4x those of call 2 in the shared scratch
    loud = wave * 4.0
    out = torch.empty_like(wave)
    binding.peak_normalize_device_noalloc(0, loud.data_ptr(), out.data_ptr(), offs.data_ptr(), n_utts, stream)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want_a)
    assert np.array_equal(loud.cpu().numpy(), (wave * 4.0).cpu().numpy())            # the input is not written to

    # call 2, straight after, in place, on the quieter original: stale maxima would show as a result 4x too small
    work = wave.clone()
    binding.peak_normalize_device_noalloc(0, work.data_ptr(), work.data_ptr(), offs.data_ptr(), n_utts, stream)
    torch.cuda.synchronize()
    assert np.array_equal(work.cpu().numpy(), want_a)

    # call 3: another batch with another utterance count (more slots than the calls before it touched)
    b = _batch(312, [5, 41, 1, 23, 8], [1, 1, 0, 0, 1])
    want_b = codec.decode_wave_batch(b[0], b[1], b[2], b[3], peak_normalize=True)
    wave_b, offs_b = _resident_unnormalised(codec, b)
    out_b = torch.empty_like(wave_b)
    binding.peak_normalize_device_noalloc(0, wave_b.data_ptr(), out_b.data_ptr(), offs_b.data_ptr(), len(b[1]) - 1, stream)
    torch.cuda.synchronize()
    assert np.array_equal(out_b.cpu().numpy(), want_b)

    # a silent utterance (peak <= 1e-8) is passed through unchanged, as the reference does on the host
    z = torch.zeros(1000, dtype=torch.float32, device="cuda:0")
    zo = torch.full_like(z, 7.0)
    zoffs = torch.tensor([0, 400, 1000], dtype=torch.int64, device="cuda:0")
    binding.peak_normalize_device_noalloc(0, z.data_ptr(), zo.data_ptr(), zoffs.data_ptr(), 2, stream)
    torch.cuda.synchronize()
    assert not zo.any().item()


def test_noalloc_refuses_more_utterances_than_its_scratch_holds(codec):
    """The scratch is a fixed 4096-slot array: a larger batch is an argument error with a message that names the limit, and is
    refused before anything is enqueued (the pointers are never touched, so any non-null value does here)."""
    import torch
    from miocodec_b200 import binding
    buf = torch.zeros(16, dtype=torch.float32, device="cuda:0")
    with pytest.raises(binding.MioError) as ei:
        binding.peak_normalize_device_noalloc(0, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), 4097)
    assert ei.value.code == binding.MIO_ERR_ARG and "MIO_NOALLOC_MAX_UTTS" in str(ei.value)
    with pytest.raises(binding.MioError):
        binding.peak_normalize_device_noalloc(0, 0, buf.data_ptr(), buf.data_ptr(), 1)             # null input
    binding.peak_normalize_device_noalloc(0, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), 0)    # an empty batch is fine
    torch.cuda.synchronize()
