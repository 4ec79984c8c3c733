"""CPU oracle for the Wheatley graph-attention hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``wheatley_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do.

Parity status
-------------
* Mask head (a6/a7), MLP (a8), PSP edge embedder (a9): pinned.  The reference
  modules import in the build container and the golden vectors under
  ``tests/golden/`` were produced by running them verbatim
  (``tests/golden/make_golden.py``).
* EGAT / EEGAT layer (a1) and GATv3 layer (a3): pinned *modulo third-party
  primitives*.  ``generic/eegatconv.py`` and ``generic/gatv3_conv.py`` were
  executed line by line in the build container on top of tiny stand-ins for
  the DGL / PyG primitives they call (``oracle/shims``: ``u_add_v``,
  ``edge_softmax``, ``u_mul_e``+``sum``; ``MessagePassing.propagate``,
  ``softmax``, ``add/remove_self_loops``).  DGL and PyG themselves are absent
  from the image (un-vendored, un-pinned dependencies), so the primitives'
  published definitions are restated in the shims.
* PyG ``GATv2Conv`` (a2): **parity unpinned** - third-party layer, not in the
  reference tree, no reference test pins it.  Restated from the GATv2
  definition and the in-repo PyG-style skeleton ``generic/gatv3_conv.py:122-198``.
"""
