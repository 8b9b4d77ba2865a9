"""Federated-learning engine: client selection, round schedule, batched local training, evaluation, aggregation."""
