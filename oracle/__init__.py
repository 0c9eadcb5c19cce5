"""oracle/ — CPU restatement of the reference's algorithms for the S2T speech-encoder hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``s2t_b200/`` may import this package; the only legal users
are ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` (where it is the thing being compared against or timed as the CPU baseline, never the
product path).

Parity pinning:
  * fbank / CMVN  — restates ``torchaudio.compliance.kaldi.fbank`` (third-party, unpinned in the
    reference's setup.py:206; torchaudio 2.11.0 on this image) as called from
    ``fairseq/data/audio/audio_utils.py:59-79`` and the reference's own ``UtteranceCMVN`` /
    ``GlobalCMVN``.  Pinned against torchaudio itself and against the reference's CMVN classes run
    from /root/reference (``oracle/make_golden.py`` → ``tests/golden/fbank_*.npz``).
  * encoder / criterion — restates the reference's PyTorch modules; pinned against the reference's
    own files executed through the import shim in ``oracle/ref_shim.py`` (``tests/golden/enc_*.npz``).
  * CTC loss — restates the alpha/beta recursion of ``torch.nn.CTCLoss`` (third-party ATen kernel the
    reference calls at ``fairseq/criterions/ctc.py:243-245,467-472``); pinned against torch 2.11 CPU.
The reference repository has no golden vectors or tests of its own for this path.
"""
