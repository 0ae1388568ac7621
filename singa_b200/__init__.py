"""singa_b200 — a Blackwell (sm_100a) native parameter-server deep-learning framework with the job API of
early SINGA (protobuf NeuralNet / Cluster / Topology configs, worker groups, server groups)."""
__version__ = "0.1.0"
