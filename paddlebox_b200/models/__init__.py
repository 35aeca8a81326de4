"""Model families: CTR-DNN, Wide&Deep, DeepFM, MMoE."""
