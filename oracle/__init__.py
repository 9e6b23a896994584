"""CPU oracle for the artspeech phoneme->articulation hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and only as the checker or as the
timed CPU baseline.  The product package ``artspeech_b200`` never imports it
and raises if its CUDA library is missing.

Contents (each function cites the reference file:line it restates; paths are
relative to the reference checkout of vribeiro1/artspeech):

* ``oracle.model``     - ArtSpeech forward restated over a plain ``state_dict``
                         (stock torch CPU ops; fp32 or fp64) and a from-the-
                         equations GRU used to cross-check ``nn.GRU``.
                         Also ``simple_artspeech_forward`` (SimpleArtSpeech) and, through
                         ``artspeech_forward.__wrapped__``, the gradient truth for the
                         training-step test.
* ``oracle.geometry``  - P2CP, Euclidean, Pearson, tract variables and ``critical_loss``
                         (CriticalLoss of principal_components/losses.py).
* ``oracle.area``      - NumPy fp64 restatement of ``area_function.py``;
  ``oracle/area_c.c``    the same chain in plain C for batches (gcc, built by
  ``oracle.build_c``     ``oracle.build_c`` into oracle/_build/).  Grid, mid_point,
                         argmatrix and area_function are PINNED bit-exactly to
                         the reference's own functions (tests/golden/area.npz);
                         intersect_semipolar_grid and evenly_spaced_fx are
                         PARITY UNPINNED: their arithmetic is inside shapely /
                         GEOS and vt_tools, absent here, and the module has no
                         caller, test or fixture in the reference.
* ``oracle.reference_loader`` / ``oracle.make_golden`` - import the real
                         reference by file path (only where /root/reference
                         exists) to pin the restatements and to write
                         ``tests/golden/*.npz``.

Where a reference module cannot be imported because of un-vendored imports
(vt_tools, vt_shape_gen, helpers chains), ``reference_loader`` compiles ONLY the
needed class / function statement, source text unmodified (``ast``), in a
namespace of the reference's own importable pieces: CriticalLoss, the
``tract_variables`` frame driver, ``save_outputs`` and the whole ``run_test``
were run that way to write tests/golden/{critical_loss.npz, tv_driver_*.csv,
run_test_info.json, model_grad.npz}.

Parity status: model / P2CP / TV restatements are pinned against outputs of
the unmodified reference run in the build container (tests/golden, script
committed).  Area function: grid / area_function pinned, the two GEOS-backed functions
unpinned (see above).
"""
