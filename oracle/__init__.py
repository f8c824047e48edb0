"""TEST INFRASTRUCTURE ONLY — CPU restatement ("oracle") of the reference's hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and there only as the checker / the timed CPU baseline.
The product package (``snn_for_object_detection_b200``) never imports it and has no
CPU fallback: it raises if the CUDA extension is missing.

PARITY UNPINNED for the neuron arithmetic: the LIF/LI update lives in the third-party
package ``norse==1.1.0`` (reference ``environment.yml:222``), which is neither vendored
in ``/root/reference`` nor installable here (no network).  ``norse_restated.py`` restates
its published algorithm; the in-tree corroboration is the author's own
``models/modules/sli.py:110-126`` which follows the same jump -> integrate -> decay order.
Everything else (event decode/binning, generator graph, anchors, box decode/NMS) IS pinned:
``tests/golden/`` holds outputs of the reference's own Python run in the build container
(``oracle/make_golden.py``), and the oracle is checked against them bit-for-bit.
"""
