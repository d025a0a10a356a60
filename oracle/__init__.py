"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT.

CPU restatement of the reference's MeshGraphNet processor hot path
(graphphysics/models/layers.py, processors.py) and of its integer graph set-up
(utils/torch_graph.py, dataset/preprocessing.py, utils/meshmask.py).

Only tests/, `__graft_entry__.smoke()` and bench.py's `cpu_baseline` /
`--impl reference` legs may import anything from here, and only as the checker
or the timed CPU baseline.  graph_physics_b200/ never imports it and raises if
its CUDA library is missing -- there is no CPU fallback.

Parity: integer side pinned by the reference's golden counts and known-answer
tensors; floating-point side is unpinned by the reference's own tests and pinned
here by fixtures produced from the reference's own source run in the build
container (tests/golden/make_golden.py).  See the module headers.
"""
