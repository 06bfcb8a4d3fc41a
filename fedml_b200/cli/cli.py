"""``fedml`` command line (reference: cli/cli.py:17-381, console script ``fedml=fedml.cli.cli:cli``):
``version``, ``login <userid> [-v release|test|dev|local] [-d]``, ``logout``, ``build -t client|server -sf -ep -cf -df``.

``login`` starts the local edge agent (cli/edge_deployment/login.py:31-440 in the reference): it binds the
device to an account id, subscribes to ``flserver_agent/<edge>/start_train`` / ``stop_train`` on the broker, and
on a start message materialises the package, rewrites its ``fedml_config.yaml`` with ``${FEDSYS.*}`` values and
spawns ``python <entry> --cf <conf> --rank <r>``, reporting the status FSM
IDLE -> INITIALIZING -> TRAINING -> STOPPING -> FINISHED.  The reference talks to open.fedml.ai; offline the
agent talks to the broker configured by ``--broker host:port`` (``LocalBroker`` by default).
``build`` zips user code into the MLOps package skeleton (``<dest>/dist-packages/{client,server}-package.zip``)."""
import json
import os
import shutil
import signal
import subprocess
import sys
import time
import zipfile

import click
import yaml

from ..arguments import substitute_fedsys_vars

HOME = os.path.join(os.path.expanduser("~"), ".fedml_b200")
PID_FILE = os.path.join(HOME, "edge-process.id")


@click.group()
def cli():
    pass


@cli.command("version", help="Display fedml version.")
def version():
    from .. import __version__
    click.echo("fedml version: " + str(__version__))


def _kill_old_agent():
    if os.path.exists(PID_FILE):
        try:
            with open(PID_FILE) as fh:
                pid = int(yaml.safe_load(fh)["process_id"])
            os.killpg(os.getpgid(pid), signal.SIGTERM)
        except Exception:  # noqa: BLE001 - stale pid file
            pass
        os.remove(PID_FILE)


@cli.command("login", help="Login to the MLOps platform and start the edge agent.")
@click.argument("userid", nargs=-1)
@click.option("--version", "-v", type=str, default="release")
@click.option("--docker", "-d", default=None, is_flag=True)
@click.option("--broker", type=str, default="127.0.0.1:1883", help="host:port of the MQTT-style broker")
@click.option("--foreground", is_flag=True, default=False)
def login(userid, version, docker, broker, foreground):
    account_id = userid[0] if userid else "0"
    os.makedirs(HOME, exist_ok=True)
    _kill_old_agent()
    if docker:                                                # the agent inside a container (reference login_with_docker.py)
        from .edge_deployment import login_with_docker_mode
        cmd = login_with_docker_mode(account_id, version)
        click.echo("docker mode: " + " ".join(cmd) + ("" if shutil.which("docker") else "   (docker is not installed: command not run)"))
        return 0
    cmd = [sys.executable, "-m", "fedml_b200.cli.edge_agent", "--account", str(account_id), "--version", version,
           "--broker", broker]
    if foreground:
        return subprocess.call(cmd)
    proc = subprocess.Popen(cmd, start_new_session=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    with open(PID_FILE, "w") as fh:
        yaml.safe_dump({"process_id": proc.pid, "account_id": account_id, "started": time.time()}, fh)
    click.echo(f"login successfully: edge agent pid {proc.pid} bound to account {account_id} ({version})")


@cli.command("logout", help="Logout from the MLOps platform and stop the edge agent.")
def logout():
    _kill_old_agent()
    click.echo("logout successfully")


@cli.command("build", help="Build packages for the MLOps platform.")
@click.option("--type", "-t", "ptype", type=click.Choice(["client", "server"]), required=True)
@click.option("--source_folder", "-sf", type=str, required=True)
@click.option("--entry_point", "-ep", type=str, required=True)
@click.option("--config_folder", "-cf", type=str, required=True)
@click.option("--dest_folder", "-df", type=str, required=True)
def build(ptype, source_folder, entry_point, config_folder, dest_folder):
    path = build_package(ptype, source_folder, entry_point, config_folder, dest_folder)
    click.echo(f"You have finished all building process.\nNow you may use {path} to start your federated learning run.")


def build_package(ptype, source_folder, entry_point, config_folder, dest_folder):
    """Layout of the reference skeleton (cli/build-package/mlops-core/fedml-{client,server}/<t>-package/):
    ``conf/fedml.yaml`` (entry + conf pointers with ${FEDSYS.*} placeholders), ``fedml/<user code>``,
    ``fedml/config/<user config>``."""
    if not os.path.isfile(os.path.join(source_folder, entry_point)):
        raise click.ClickException(f"entry point {entry_point} not found in {source_folder}")
    stage = os.path.join(dest_folder, f".{ptype}-package-stage")
    shutil.rmtree(stage, ignore_errors=True)
    pkg = os.path.join(stage, f"{ptype}-package")
    os.makedirs(os.path.join(pkg, "conf"))
    shutil.copytree(source_folder, os.path.join(pkg, "fedml"), ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    shutil.copytree(config_folder, os.path.join(pkg, "fedml", "config"), dirs_exist_ok=True)
    conf = {"entry_config": {"entry_file": os.path.join("fedml", entry_point), "conf_file": "config/fedml_config.yaml"},
            "dynamic_args": {"rank": "${FEDSYS.CLIENT_INDEX}" if ptype == "client" else "0", "run_id": "${FEDSYS.RUN_ID}",
                             "client_id_list": "${FEDSYS.CLIENT_ID_LIST}", "client_num_in_total": "${FEDSYS.CLIENT_NUM}",
                             "mqtt_config_path": "${FEDSYS.MQTT_CONFIG_PATH}", "s3_config_path": "${FEDSYS.S3_CONFIG_PATH}",
                             "log_file_dir": "${FEDSYS.LOG_FILE_DIR}"}}
    with open(os.path.join(pkg, "conf", "fedml.yaml"), "w") as fh:
        yaml.safe_dump(conf, fh)
    out_dir = os.path.join(dest_folder, "dist-packages")
    os.makedirs(out_dir, exist_ok=True)
    zpath = os.path.join(out_dir, f"{ptype}-package.zip")
    with zipfile.ZipFile(zpath, "w", zipfile.ZIP_DEFLATED) as z:
        for root, _, files in os.walk(pkg):
            for f in files:
                full = os.path.join(root, f)
                z.write(full, os.path.relpath(full, stage))
    shutil.rmtree(stage, ignore_errors=True)
    return zpath


def materialise_package(zip_path, work_dir, fedsys: dict):
    """Edge-agent side of ``build``: unzip, substitute ``${FEDSYS.*}`` into the package's fedml_config.yaml, return
    (entry_file, conf_file, dynamic_args)."""
    shutil.rmtree(work_dir, ignore_errors=True)
    with zipfile.ZipFile(zip_path) as z:
        z.extractall(work_dir)
    pkg = os.path.join(work_dir, [d for d in os.listdir(work_dir) if d.endswith("-package")][0])
    with open(os.path.join(pkg, "conf", "fedml.yaml")) as fh:
        conf = yaml.safe_load(substitute_fedsys_vars(fh.read(), fedsys))
    entry = os.path.join(pkg, conf["entry_config"]["entry_file"])
    conf_file = os.path.join(pkg, "fedml", conf["entry_config"]["conf_file"])
    with open(conf_file) as fh:
        user_cfg = yaml.safe_load(substitute_fedsys_vars(fh.read(), fedsys)) or {}
    dyn = conf.get("dynamic_args", {})
    for fam in user_cfg.values():
        if isinstance(fam, dict):
            for k in list(fam):
                if k in dyn and k != "rank":
                    fam[k] = dyn[k]
    with open(conf_file, "w") as fh:
        yaml.safe_dump(user_cfg, fh)
    return entry, conf_file, dyn


if __name__ == "__main__":
    cli()
