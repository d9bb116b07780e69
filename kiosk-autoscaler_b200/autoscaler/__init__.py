"""kiosk-autoscaler, B200-native tally/decision path: same package surface as
the reference (``autoscaler.Autoscaler``, ``autoscaler.redis.RedisClient``;
reference autoscaler/__init__.py:30-32) plus the batched replay modules."""
from autoscaler import redis
from autoscaler import engine
from autoscaler import snapshot
from autoscaler import replay
from autoscaler.autoscaler import Autoscaler

__all__ = ['Autoscaler', 'redis', 'engine', 'snapshot', 'replay']
