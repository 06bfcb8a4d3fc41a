"""Beehive server: aggregates the ``.mnn`` model files phones upload (read / averaged / written back on the FlatBuffers
bytes, no MNN runtime needed) and publishes the new global file through MQTT + the object store."""
import fedml

if __name__ == "__main__":
    fedml.run_mnn_server()
