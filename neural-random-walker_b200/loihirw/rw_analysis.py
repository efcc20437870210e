"""Probe rasters -> walker counts.  Same function names, arguments and return values as the
reference's loihirw/rw_analysis.py, without its ``from nxsdk... import *`` (rw_analysis.py:8), and
vectorised (prefix sums instead of a Python loop per bin).

A "probe" is any object with ``.data`` indexable as ``data[neuron][time]`` (NxSDK spike/voltage
probes in the reference; ``loihirw.probes.Probe`` here).
"""
import logging
import sys

import numpy as np

rw_logger = logging.getLogger("NRW")


def _row(probe, neuron):
    return np.asarray(probe.data[:][neuron])


def _three_digit_counter(probe):
    """Counter neurons hold a base-1000 number in three compartments (rw_analysis.py:23,28)."""
    d = probe.data[:]
    return d[0][0] + 1000 * d[1][0] + 1000 * 1000 * d[2][0]


def _unwrap_titles(track_probes):
    """[title, probe] pairs -> probe, in place (rw_analysis.py:74-77; :43-46 in count_steps)."""
    for k in range(len(track_probes)):
        if track_probes[k] and isinstance(track_probes[k][0], str):
            track_probes[k] = track_probes[k][1]


def _segment_sums(boundary_raster, spike_raster):
    """Sums of ``spike_raster`` over the bins delimited by the non-zeros of ``boundary_raster``:
    bin i = [previous boundary + 1, boundary_i)  (rw_analysis.py:90-99, :168-178)."""
    ends = np.nonzero(boundary_raster)[0]
    if len(ends) == 0:
        return np.zeros(0), ends
    starts = np.concatenate([[0], ends[:-1] + 1])
    cs = np.concatenate([[0], np.cumsum(np.asarray(spike_raster, dtype=np.float64))])
    return cs[ends] - cs[np.minimum(starts, ends)], ends


def process_outputs_ends(track_probes, probe_start=4, probe_stride=5, num_meshpoints=41, num_runs=1):
    """Track 1/3: final counter voltages = cumulative density per mesh point (rw_analysis.py:11-36).
    Returns (count[num_meshpoints, num_runs], num_steps[1, num_runs])."""
    count = np.zeros((num_meshpoints, num_runs))
    num_steps = np.zeros((1, num_runs))
    runs = [track_probes] if num_runs == 1 else [track_probes[k] for k in range(num_runs)]
    for k, tp in enumerate(runs):
        num_steps[0, k] = _three_digit_counter(tp[1][0]) / 128.
        for j in range(num_meshpoints):
            count[j, k] = _three_digit_counter(tp[probe_start + j * probe_stride][0])
    return count, num_steps


def count_steps(track_probes, probe_start=4, probe_stride=5, num_meshpoints=41, num_runs=1):
    """Model steps a Track 1/3 run lasted, from the supervisor counter (rw_analysis.py:38-55)."""
    for k in range(num_runs):
        if track_probes[k] and isinstance(track_probes[k][0], str):
            track_probes[k] = track_probes[k][1]
    num_steps = np.zeros((num_runs, 1))
    if num_runs == 1:
        num_steps[0] = _three_digit_counter(track_probes[1][0]) / 128.
    else:
        for k in range(num_runs):
            num_steps[k] = _three_digit_counter(track_probes[k][1][0]) / 128.
    return num_steps


def process_outputs(track_probes, probe_start=3, probe_stride=5, num_meshpoints=41):
    """Track 2/5 (flat probe list, ``probe_stride`` probes per mesh point) -> density over time
    (rw_analysis.py:57-101).  Returns (count[num_meshpoints, num_bins + 1], num_bins)."""
    _unwrap_titles(track_probes)
    if num_meshpoints <= 0:
        return 0, int(np.size(np.nonzero(_row(track_probes[0][0], 0))))
    num_bins = int(np.size(np.nonzero(_row(track_probes[probe_start][0], 0))))
    count = np.zeros((num_meshpoints, num_bins + 1))
    for j in range(num_meshpoints):
        probe = track_probes[probe_start + j * probe_stride][0]
        sums, ends = _segment_sums(_row(probe, 0), _row(probe, 1))
        num_bins = len(ends)
        count[j, :num_bins] = sums
    return count, num_bins


def process_outputs_verbose(track_probes, probe_start=3, probe_stride=5, num_meshpoints=41):
    """process_outputs plus the reference's debugging log lines (rw_analysis.py:103-140)."""
    count, num_bins = process_outputs(list(track_probes), probe_start, probe_stride, num_meshpoints)
    for j in range(num_meshpoints):
        probe = track_probes[probe_start + j * probe_stride][0]
        logging.info('mesh_node {}: {}'.format(j, np.nonzero(_row(probe, 0))))
    np.set_printoptions(threshold=sys.maxsize)
    return count, num_bins


def process_outputs_split(track_probes, probe_start=0, probe_stride=1, num_meshpoints=41):
    """Track 6 (supervisors and mesh nodes in separate groups) -> density over time
    (rw_analysis.py:142-179); the decoder both miniapps use (fluence_mini_app.py:163).

    ``track_probes[0]`` = supervisor groups, ``track_probes[1][j]`` = ``[title, probe]`` of mesh
    point j whose probe rows are (0) the model-step boundary raster and (1) one spike per walker.
    Returns (count[num_meshpoints, num_bins + 1], num_bins)."""
    super_probes, mesh_probes = track_probes[0], track_probes[1]
    num_bins = int(np.size(np.nonzero(_row(super_probes[0][1], 0))))
    count = np.zeros((num_meshpoints, num_bins + 1))
    for j in range(num_meshpoints):
        probe = mesh_probes[probe_start + j * probe_stride][1]
        sums, ends = _segment_sums(_row(probe, 0), _row(probe, 1))
        num_bins = len(ends)
        count[j, :num_bins] = sums
    return count, num_bins
