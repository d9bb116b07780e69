"""Turn on and off k8s resources based on items in the Redis queue.

Same environment variables, defaults and casts as the reference entry point
(scale.py:74-92 of the reference): REDIS_HOST, REDIS_PORT, REDIS_INTERVAL,
QUEUES, QUEUE_DELIMITER, INTERVAL, RESOURCE_NAMESPACE, RESOURCE_TYPE,
RESOURCE_NAME (required), MIN_PODS, MAX_PODS, KEYS_PER_POD.
"""
import gc
import logging
import logging.handlers
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import autoscaler  # noqa: E402

_MISSING = object()


def config(name, default=_MISSING, cast=str):
    """``decouple.config`` when python-decouple is installed, else the same
    lookup over os.environ (value cast with ``cast``; a missing variable with no
    default is an error, as for RESOURCE_NAME in the reference)."""
    try:
        import decouple
        if default is _MISSING:
            return decouple.config(name, cast=cast)
        return decouple.config(name, default=default, cast=cast)
    except ImportError:
        pass
    if name in os.environ:
        return cast(os.environ[name])
    if default is _MISSING:
        raise KeyError('{} not found. Declare it as envvar or define a default value.'.format(name))
    return cast(default)


def load_config():
    """All settings of the reference's scale.py, by the same names."""
    return dict(
        redis_host=config('REDIS_HOST', cast=str, default='redis-master'),
        redis_port=config('REDIS_PORT', default=6379, cast=int),
        redis_backoff=config('REDIS_INTERVAL', default=1, cast=int),
        queues=config('QUEUES', default='predict,track', cast=str),
        queue_delim=config('QUEUE_DELIMITER', ',', cast=str),
        interval=config('INTERVAL', default=5, cast=int),
        namespace=config('RESOURCE_NAMESPACE', default='default'),
        resource_type=config('RESOURCE_TYPE', default='deployment'),
        name=config('RESOURCE_NAME'),
        min_pods=config('MIN_PODS', default=0, cast=int),
        max_pods=config('MAX_PODS', default=1, cast=int),
        keys_per_pod=config('KEYS_PER_POD', default=1, cast=int))


def initialize_logger(debug_mode=True):
    logger = logging.getLogger()
    logger.setLevel(logging.DEBUG)
    formatter = logging.Formatter('[%(asctime)s]:[%(levelname)s]:[%(name)s]: %(message)s')
    console = logging.StreamHandler(stream=sys.stdout)
    console.setFormatter(formatter)
    console.setLevel(logging.DEBUG if debug_mode else logging.INFO)
    fh = logging.handlers.RotatingFileHandler(filename='autoscaler.log', maxBytes=10000000,
                                              backupCount=10)
    fh.setFormatter(formatter)
    fh.setLevel(logging.DEBUG)
    logger.addHandler(console)
    logger.addHandler(fh)
    logging.getLogger('kubernetes.client.rest').setLevel(logging.INFO)


if __name__ == '__main__':
    initialize_logger()
    _logger = logging.getLogger(__file__)
    CFG = load_config()
    REDIS_CLIENT = autoscaler.redis.RedisClient(host=CFG['redis_host'], port=CFG['redis_port'],
                                                backoff=CFG['redis_backoff'])
    SCALER = autoscaler.Autoscaler(redis_client=REDIS_CLIENT, queues=CFG['queues'],
                                   queue_delim=CFG['queue_delim'])
    while True:
        try:
            SCALER.scale(namespace=CFG['namespace'], resource_type=CFG['resource_type'],
                         name=CFG['name'], min_pods=CFG['min_pods'], max_pods=CFG['max_pods'],
                         keys_per_pod=CFG['keys_per_pod'])
            gc.collect()
            time.sleep(CFG['interval'])
        except Exception as err:  # pylint: disable=broad-except
            _logger.critical('Fatal Error: %s: %s', type(err).__name__, err)
            sys.exit(1)
