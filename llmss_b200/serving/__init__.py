"""Serving front ends: broker (RESP list queue), HTTP producer, gRPC service, TP consumer loop."""
