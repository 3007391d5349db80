"""CPU oracle for the NIC-ES / NIC-NES population-evaluation hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(`nes-img-captioning_b200/`) may import this package; only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py` do, and there only as the checker / the timed CPU arm.

What is pinned against what
---------------------------
* fc_caption decode, member construction, centred ranks, gradient estimate,
  SGD and Adam are restatements of the reference's Python
  (rubencart/NES-img-captioning, /root/reference) and are PINNED: they were
  checked bit-for-bit against the imported reference in the build container by
  `oracle/refcheck.py`, and `tests/golden/make_golden.py` (which imports the
  reference) generated the committed fixtures under `tests/golden/`.
* The shared noise table, its padded row layout and the offset sampler do not
  exist in the reference (it draws an unseeded torch normal per worker, see
  src/algorithm/nets.py:92-102).  They are this build's own specification,
  written down in `oracle/noise.py`; for those three items PARITY IS UNPINNED
  against the reference and pinned only against Philox4x32-10's published
  known-answer vectors plus this restatement.
"""
