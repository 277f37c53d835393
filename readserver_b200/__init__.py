"""B200-native engine for ReadServer's FM-index hot path (count + locate over run-length BWTs)."""
__version__ = "0.1.0"
