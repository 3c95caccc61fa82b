"""CPU oracle for the EEND-EDA hot path.  TEST INFRASTRUCTURE ONLY.

Everything under ``oracle/`` is a CPU restatement (torch fp32 / numpy fp64) of the
reference's algorithm for the path BASELINE.json:north_star names.  It exists to
CHECK the CUDA path; it is never the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it.  Nothing in ``speaker_diarization_b200/`` imports it, and the
product path raises if the CUDA library is missing.

Generators of the committed fixtures (build container only; they import /root/reference):
``make_golden.py`` (small model / PIT / front-end cases), ``make_golden_full.py`` (configs[1] and configs[4] at full
size; the reference's own driver ``infer_eda.infer`` run unmodified on 16 kHz sessions and on long recordings with 1-6
speakers), ``make_golden_train.py``, ``make_golden_conformer.py``, ``make_golden_eend.py``, ``make_golden_labels.py``,
``make_golden_transforms.py`` (feature.transform None / 'log' / 'logmel23_espnet'), ``make_golden_rttm.py`` (the
unmodified ``bin/make_rttm.py`` via runpy + the vendored md-eval.pl).  ``build_ref.py`` places the reference's own
three files of the path into the git-ignored ``oracle/_ref`` for the timed reference arm of bench.py.

Parity status
-------------
* model path (``eend_eda_oracle``): PINNED.  Checked against the reference's own
  ``speaker_diarization/eend_eda/models.py`` imported unchanged in the build
  container; the outputs are frozen under ``tests/golden/`` by
  ``oracle/make_golden.py`` and re-checked by ``tests/test_oracle_golden.py``.
* feature path (``feature_oracle``): the reference's ``feature.py`` needs
  ``librosa==0.10.2`` (requirements.txt:15), which is not vendored and cannot be
  installed.  ``librosa.stft`` / ``librosa.filters.mel`` are restated from their
  published algorithm (``oracle/librosa_stub.py``) and cross-checked against an
  independent second statement (``torch.stft`` + ``torchaudio`` Slaney filterbank;
  ``tests/test_feature_oracle.py``).
  The reference's ``feature.py`` itself then runs unmodified on top of the stub and
  its outputs are frozen as goldens.  With respect to librosa proper this leg is
  "parity unpinned" (no librosa to execute); with respect to ``feature.py`` it is
  pinned.
"""
