"""``densityrw`` on a B200: the public surface of the reference's density-random-walk package
(``load_transitions``, ``Transition``, ``DensityRW``, ``run_miniapp``, the ring test graph), backed by
the GPU walk engine instead of the Fugu spiking simulator."""
