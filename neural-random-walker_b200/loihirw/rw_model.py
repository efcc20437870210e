"""``rw_model`` on a B200: the runner entry points of the reference's loihirw/rw_model.py
(``test_network_short`` :20-68, ``rad_test`` :70-93) with the same signatures and return tuple, on
the GPU walk engine instead of an NxSDK board."""
import logging

from .rw_mesh import makeNetwork

rw_logger = logging.getLogger("NRW")


def test_network_short(custom_network, x, y, time, core_density=10, location=[(0, 0)],
                       num_walkers=100, track=0, verbose=0, input_times=[1], **walk_args):
    """rw_model.py:20-68.  Returns ``(tprobe1, eprobe1, track_probes, core_neuroncount, max_coreid)``;
    the Loihi execution-time / energy probes (``track == 2``, :53-57) have no GPU counterpart and
    are returned as empty lists, as the reference does for every other track."""
    rw_logger.info('Setting up network')
    core_neuroncount, max_coreid, track_probes, chip_count, _ = makeNetwork(
        superNet=None, x=x, y=y, custom_network=custom_network, core_density=core_density,
        input_x=location, sim_time=time, num_walkers=num_walkers, track=track, verbose=verbose,
        input_times=input_times, s=None, **walk_args)
    rw_logger.info("Ran on 1 GPU (Loihi placement would be {} cores on {} chips)".format(
        max_coreid, chip_count))
    return [], [], track_probes, core_neuroncount, max_coreid


test_network_short.__test__ = False        # a runner, not a pytest test


def rad_test(custom_network, location, time=1000, num_walkers=100, verbose=0, core_density=10,
             input_times=[1], track=1, size=(901, 1), **walk_args):
    """rw_model.py:70-93: walk ``num_walkers`` walkers from every mesh point in ``location`` for
    ``time`` model steps on the mesh stored in ``custom_network``.  Additive keyword arguments
    ``rng`` ('auto' | 'replay' | 'philox'), ``sampler`` and ``spike_mode`` are passed to the walk."""
    return test_network_short(custom_network, size[0], size[1], time, core_density, location,
                              num_walkers, track, verbose=verbose, input_times=input_times,
                              **walk_args)
