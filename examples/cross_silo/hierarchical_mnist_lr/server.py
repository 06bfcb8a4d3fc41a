import fedml

if __name__ == "__main__":
    fedml.run_hierarchical_cross_silo_server()
