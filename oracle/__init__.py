"""CPU oracle for the Point-NeRF per-ray hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in ``pointnerf_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker.

Parity status: the reference (yjcaimeow/pointnerf) ships no tests, golden
vectors or fixtures for this path, so parity is pinned
against OUTPUTS OF THE REFERENCE ITSELF produced in the build container:

* ``PointAggregator`` / ``ray_march`` / ``near_far_linear_ray_generation`` /
  ``positional_encoding`` are imported unmodified from ``/root/reference``
  (``oracle/ref_import.py``) and their outputs + autograd gradients on seeded
  inputs are committed under ``tests/golden/`` by ``oracle/make_golden.py``.
* the six pycuda kernels of ``lighting_fast_querier`` are compiled, from the
  source string where it lies in the reference tree, as serial host C++
  (``oracle/build_ref.py`` -> ``oracle/_ref/libpn_ref_querier.so``) and run in
  ascending-thread order (the canonical schedule).
* the four kernels of the second, perspective-frustum ``lighting_fast_querier``
  (``models/neural_points/query_point_indices.py``) the same way ->
  ``oracle/_ref/libpn_ref_querier_frustum.so``, driven by ``oracle/frustum_ref.py``;
  golden vectors ``tests/golden/frustum_*_ref.npz``; ``oracle/frustum_model.py`` is
  a numpy statement of the product's CSR formulation, checked against them.
* ``oracle/ref_fragments.py`` compiles whole classes / single functions / one ``if``
  block of reference files that cannot be imported (pycuda, cv2, torch_scatter at
  module scope) from their source where it lies: both ``lighting_fast_querier``
  classes run end to end on the CPU (own glue + own kernels), and the maintenance
  steps (prune, grow, probe block, unmask, voxel down-sampling) are pinned this way.
* ``oracle/querier_port.c`` and ``oracle/torch_port.py`` are the restatements
  that travel to the GPU box; they are checked against the two items above.
"""
