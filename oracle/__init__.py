"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32) of the ZipVoice hot path.

Nothing under ``oracle/`` is part of the shipped product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker / the timed CPU baseline.

Parity status: PINNED for the model path (text encoder, fm_decoder, CFG wrapper,
Euler solver, text-condition helpers, dialog speaker embedding) and for the
VocosFbank mel front-end: ``oracle/make_golden.py`` imports the unmodified
reference from /root/reference in the build container (``lhotse`` / ``librosa`` replaced
by the stand-ins in ``oracle/shims``), feeds it the weights of
``zipvoice_b200.synth.make_state_dict`` and stores its outputs under ``tests/golden``;
``tests/test_oracle_golden.py`` checks this restatement against those files.
Parity UNPINNED for (i) the Vocos vocoder (third-party ``vocos`` package, not vendored
in the reference, no weights, no tests): ``oracle/vocos_oracle.py`` restates the
published architecture and is its own oracle; (ii) the one-frame boundary of lhotse's
``compute_num_frames`` (source absent, restated as round-half-up in ``fbank_oracle``).
The reference is Python: there is no compiled ``oracle/_ref``.
"""
