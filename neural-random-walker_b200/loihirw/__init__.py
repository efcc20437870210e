"""``loihirw`` on a B200: the module and function names of the reference's Loihi package
(loihirw/rw_mesh.py, rw_model.py, rw_core.py, rw_analysis.py), re-targeted at the GPU walk engine.

The reference builds spiking circuits with Intel's proprietary NxSDK and reads spike probes back;
here ``rw_mesh`` turns the transition-matrix file into device CSR tables, ``rw_model.rad_test`` runs
the walk with libnrw and returns probe-shaped rasters, and ``rw_analysis`` (NxSDK-free, same
function names and return values) decodes them into per-node, per-model-step walker counts.
"""
