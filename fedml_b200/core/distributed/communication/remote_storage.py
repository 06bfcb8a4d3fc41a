"""Model blob store for the MQTT_S3 data plane (reference: communication/mqtt_s3/remote_storage.py:39-120).

The reference dumps the model with joblib to an S3 object and ships the key + presigned URL over MQTT;
downloads retry three times.  Offline there is no S3, so the default store is a shared directory
(``LocalObjectStore``); with ``boto3`` installed and an ``s3_config_path`` it can talk to a real bucket.  Blobs
are ``Message``-codec tensor containers (raw buffers) rather than pickles."""
import os
import time
import uuid

from .message import Message


class LocalObjectStore:
    def __init__(self, root):
        self.root = root
        os.makedirs(root, exist_ok=True)

    def _path(self, key):
        return os.path.join(self.root, key.replace("/", "_"))

    def write_model(self, key, model) -> str:
        m = Message()
        m.add_params("model", model)
        tmp = self._path(key) + f".{uuid.uuid4().hex}.tmp"
        with open(tmp, "wb") as fh:
            fh.write(m.to_bytes())
        os.replace(tmp, self._path(key))
        return "file://" + os.path.abspath(self._path(key))

    def read_model(self, key, retries=3):
        last = None
        for attempt in range(retries):
            try:
                with open(self._path(key), "rb") as fh:
                    return Message.from_bytes(fh.read()).get("model")
            except OSError as exc:
                last = exc
                time.sleep(0.05 * (attempt + 1))
        raise last

    def upload_file(self, src_path, key) -> str:
        import shutil
        shutil.copyfile(src_path, self._path(key))
        return "file://" + os.path.abspath(self._path(key))

    def download_file(self, key, dst_path):
        import shutil
        os.makedirs(os.path.dirname(os.path.abspath(dst_path)) or ".", exist_ok=True)
        shutil.copyfile(self._path(key), dst_path)
        return dst_path

    def delete(self, key) -> bool:
        try:
            os.remove(self._path(key))
            return True
        except FileNotFoundError:
            return False


class S3Storage(LocalObjectStore):
    """Real S3 when boto3 and credentials exist; same interface."""

    def __init__(self, s3_config_path):
        import boto3
        import yaml
        with open(s3_config_path) as fh:
            cfg = yaml.safe_load(fh)
        self.bucket = cfg["BUCKET_NAME"]
        self.s3 = boto3.client("s3", region_name=cfg.get("CN_REGION_NAME"), aws_access_key_id=cfg.get("CN_S3_AKI"),
                               aws_secret_access_key=cfg.get("CN_S3_SAK"))

    def write_model(self, key, model):
        m = Message(); m.add_params("model", model)
        self.s3.put_object(Body=m.to_bytes(), Bucket=self.bucket, Key=key)
        return self.s3.generate_presigned_url("get_object", ExpiresIn=60 * 60 * 24 * 5,
                                              Params={"Bucket": self.bucket, "Key": key})

    def read_model(self, key, retries=3):
        for attempt in range(retries):
            try:
                body = self.s3.get_object(Bucket=self.bucket, Key=key)["Body"].read()
                return Message.from_bytes(body).get("model")
            except Exception:  # noqa: BLE001
                if attempt == retries - 1:
                    raise
                time.sleep(1)


    def upload_file(self, src_path, key) -> str:
        self.s3.upload_file(src_path, self.bucket, key)
        return self.s3.generate_presigned_url("get_object", ExpiresIn=60 * 60 * 24 * 5, Params={"Bucket": self.bucket, "Key": key})

    def download_file(self, key, dst_path):
        os.makedirs(os.path.dirname(os.path.abspath(dst_path)) or ".", exist_ok=True)
        self.s3.download_file(self.bucket, key, dst_path)
        return dst_path

    def delete(self, key) -> bool:
        self.s3.delete_object(Bucket=self.bucket, Key=key)
        return True


def create_store(args):
    path = getattr(args, "s3_config_path", None)
    if path and os.path.exists(str(path)):
        try:
            return S3Storage(path)
        except Exception:  # noqa: BLE001 - boto3 missing or no network: fall back to the shared directory
            pass
    root = getattr(args, "object_store_dir", None) or os.path.join(getattr(args, "model_file_cache_folder", "./model_file_cache"),
                                                                   f"object_store_run_{getattr(args, 'run_id', 0)}")
    return LocalObjectStore(root)
