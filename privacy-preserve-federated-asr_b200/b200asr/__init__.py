"""b200asr: B200-native client-local ASR update + FedAvg (host side; kernels live in libb200asr.so)."""
