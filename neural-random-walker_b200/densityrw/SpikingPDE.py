"""The one type of the reference's densityrw/SpikingPDE.py that is part of the walk's input
surface: ``Transition`` (SpikingPDE.py:273-279).  The spiking simulator itself (Neuron, Spike, Unit,
MarkovNetwork, run_simulation) is out of scope -- see DESIGN.md."""

max_value = 100.0          # densityrw/SpikingConfig.py:13


class Transition:
    """Outgoing edges of one node: ``location`` is the node key (a tuple such as ``(3,)``),
    ``neighbors`` a list of ``(destination_location, probability)``."""

    def __init__(self, **kwargs):
        if 'location' in kwargs:
            self.location = kwargs['location']
        self.neighbors = kwargs['neighbors'] if 'neighbors' in kwargs else []

    def __repr__(self):
        return "Transition(location=%r, neighbors=%r)" % (getattr(self, 'location', None), self.neighbors)


def reset_neuron_list():
    """No spiking neurons are instantiated on the GPU path (SpikingPDE.py global list)."""
    return None
