"""TLS material handling (API parity with reference ``metisfl/utils/ssl_configurator.py:16-77``).

The reference ships a fixed self-signed pair under ``resources/ssl_config/default`` made by
``gen_certificates.sh``; here the default pair is generated on first use with ``cryptography``
(pyOpenSSL is not installed) and cached in the same place.
"""
import datetime
import ipaddress
import os

from metisfl_b200.proto.metis_pb2 import SSLConfig, SSLConfigFiles, SSLConfigStream
from metisfl_b200.utils import proto_messages_factory as proto_factory
from metisfl_b200.utils.metis_logger import MetisLogger


class SSLConfigurator(object):
    DIR = os.path.dirname(__file__)
    DEFAULT_DIR = os.path.join(DIR, "..", "resources", "ssl_config", "default")

    @classmethod
    def _generate_self_signed(cls, cert_path, key_path, hostnames=("localhost",), days=3650):
        from cryptography import x509
        from cryptography.hazmat.primitives import hashes, serialization
        from cryptography.hazmat.primitives.asymmetric import rsa
        from cryptography.x509.oid import NameOID
        key = rsa.generate_private_key(public_exponent=65537, key_size=2048)
        name = x509.Name([x509.NameAttribute(NameOID.COMMON_NAME, hostnames[0]),
                          x509.NameAttribute(NameOID.ORGANIZATION_NAME, "metisfl_b200")])
        alt = [x509.DNSName(h) for h in hostnames] + [x509.IPAddress(ipaddress.ip_address("127.0.0.1"))]
        now = datetime.datetime.now(datetime.timezone.utc)
        cert = (x509.CertificateBuilder().subject_name(name).issuer_name(name).public_key(key.public_key())
                .serial_number(x509.random_serial_number()).not_valid_before(now - datetime.timedelta(days=1))
                .not_valid_after(now + datetime.timedelta(days=days))
                .add_extension(x509.SubjectAlternativeName(alt), critical=False)
                .sign(key, hashes.SHA256()))
        os.makedirs(os.path.dirname(cert_path), exist_ok=True)
        with open(key_path, "wb") as f:
            f.write(key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.TraditionalOpenSSL,
                                      serialization.NoEncryption()))
        with open(cert_path, "wb") as f:
            f.write(cert.public_bytes(serialization.Encoding.PEM))

    @classmethod
    def gen_default_certificates(cls, as_stream=False):
        public_certificate = os.path.abspath(os.path.join(cls.DEFAULT_DIR, "server-cert.pem"))
        private_key = os.path.abspath(os.path.join(cls.DEFAULT_DIR, "server-key.pem"))
        if not (os.path.exists(public_certificate) and os.path.exists(private_key)):
            cls._generate_self_signed(public_certificate, private_key)
        if as_stream:
            return cls.load_file_as_stream(public_certificate), cls.load_file_as_stream(private_key)
        return public_certificate, private_key

    @classmethod
    def load_certificates_from_ssl_config_pb(cls, ssl_config_pb: SSLConfig, as_stream=False):
        public_certificate, private_key = None, None
        if ssl_config_pb is None or not ssl_config_pb.enable_ssl:
            return public_certificate, private_key
        which = ssl_config_pb.WhichOneof("config")
        if which == "ssl_config_files":
            public_certificate = ssl_config_pb.ssl_config_files.public_certificate_file
            private_key = ssl_config_pb.ssl_config_files.private_key_file
            if as_stream:
                public_certificate = cls.load_file_as_stream(public_certificate)
                private_key = cls.load_file_as_stream(private_key)
        elif which == "ssl_config_stream":
            public_certificate = ssl_config_pb.ssl_config_stream.public_certificate_stream
            private_key = ssl_config_pb.ssl_config_stream.private_key_stream
        else:
            MetisLogger.warning("SSL was requested but no certificate was provided; proceeding without SSL.")
        return public_certificate, private_key

    @classmethod
    def load_file_as_stream(cls, filepath):
        if filepath and os.path.exists(filepath):
            with open(filepath, "rb") as f:
                return f.read()
        return None

    @classmethod
    def gen_public_ssl_config_pb_as_stream(cls, ssl_config_pb: SSLConfig):
        """Copy of the config that carries only the PUBLIC certificate, as an in-message stream."""
        config_pb = None
        if ssl_config_pb.enable_ssl:
            which = ssl_config_pb.WhichOneof("config")
            stream = None
            if which == "ssl_config_files":
                stream = cls.load_file_as_stream(ssl_config_pb.ssl_config_files.public_certificate_file)
            elif which == "ssl_config_stream":
                stream = ssl_config_pb.ssl_config_stream.public_certificate_stream
            config_pb = proto_factory.MetisProtoMessages.construct_ssl_config_stream_pb(
                public_certificate_stream=stream)
        return proto_factory.MetisProtoMessages.construct_ssl_config_pb(
            enable_ssl=ssl_config_pb.enable_ssl, config_pb=config_pb)


if __name__ == "__main__":  # python -m metisfl_b200.utils.ssl_configurator --out DIR [--host NAME ...]
    import argparse
    ap = argparse.ArgumentParser(description="write a self-signed server-cert.pem / server-key.pem pair")
    ap.add_argument("--out", required=True)
    ap.add_argument("--host", action="append", default=None)
    ap.add_argument("--days", type=int, default=3650)
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    SSLConfigurator._generate_self_signed(os.path.join(a.out, "server-cert.pem"), os.path.join(a.out, "server-key.pem"),
                                          tuple(a.host or ["localhost"]), a.days)
    print("wrote", a.out)
