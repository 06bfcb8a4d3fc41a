from setuptools import find_packages, setup

setup(
    name="fedml_b200",
    version="0.7.39+b200.1",
    description="Blackwell-native federated learning framework with the FedML API surface",
    packages=find_packages(include=["fedml_b200", "fedml_b200.*", "fedml", "fedml.*"]),
    package_data={"fedml_b200": ["config/*/*.yaml", "native/*.so"]},
    python_requires=">=3.10",
    install_requires=["torch", "numpy", "pyyaml", "click"],
    entry_points={"console_scripts": ["fedml=fedml_b200.cli.cli:cli"]},
)
