"""grappa_b200: B200-native hot path of Grappa (GNN -> bonded-parameter readout -> MM energy + forces)."""
__version__ = "0.1.0"
