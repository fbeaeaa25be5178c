"""Metrics schema + validators (parity with /root/reference/albert/metrics_utils.py:1-24, pydantic v2 spelling)."""
from typing import Dict, List, Tuple

from pydantic import BaseModel, StrictFloat, confloat, conint

from ...dht import BytesWithPublicKey, RecordValidatorBase, RSASignatureValidator, SchemaValidator


class LocalMetrics(BaseModel):
    step: conint(ge=0, strict=True)
    samples_per_second: confloat(ge=0.0, strict=True)
    samples_accumulated: conint(ge=0, strict=True)
    loss: StrictFloat
    mini_steps: conint(ge=0, strict=True)

    def dict(self, *args, **kwargs):  # pydantic-v1 spelling used at run_trainer.py:163
        return self.model_dump(*args, **kwargs)

    @classmethod
    def parse_obj(cls, obj):  # pydantic-v1 spelling used at run_first_peer.py:182
        return cls.model_validate(obj)


class MetricSchema(BaseModel):
    metrics: Dict[BytesWithPublicKey, LocalMetrics]


def make_validators(experiment_prefix: str) -> Tuple[List[RecordValidatorBase], bytes]:
    signature_validator = RSASignatureValidator()
    validators = [SchemaValidator(MetricSchema, prefix=experiment_prefix), signature_validator]
    return validators, signature_validator.local_public_key
