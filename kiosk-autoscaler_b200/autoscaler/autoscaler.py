"""Drop-in ``Autoscaler`` whose tally and decision arithmetic run on a B200.

Same constructor, attributes, methods, argument meaning and error behaviour as
the reference class (autoscaler/autoscaler.py:37-273 of the reference); the
Kubernetes wrappers stay plain client calls, while ``tally_queues``,
``clip_pod_count`` and ``get_desired_pods`` go through libkas (include/kas.h).
``replay`` is new: the same ``scale`` loop run over a whole trace and a set of
policies in one kernel.
"""
from __future__ import annotations

import logging
import timeit

from autoscaler import engine
from autoscaler import snapshot as _snapshot


def _kubernetes():
    import kubernetes  # deferred: only the live-cluster wrappers need it
    return kubernetes


def _is_api_exception(err):
    """``except kubernetes.client.rest.ApiException`` of the reference
    (autoscaler.py:95,111,127,143,267) also catches subclasses; matched by name over
    the MRO so that the kubernetes package is only needed on a live cluster."""
    return any(c.__name__ == 'ApiException' for c in type(err).__mro__)


class Autoscaler(object):
    """Read Redis and scale up k8s pods if required.

    Args:
        redis_client: object exposing ``llen(name)`` and either
            ``key_snapshot()`` (offline replay) or
            ``scan_iter(match=, count=)`` (a live server).
        queues: string, queue names joined by ``queue_delim``.
        queue_delim: string, delimiter between queue names.
    """

    def __init__(self, redis_client, queues='predict', queue_delim=','):
        # duplicates collapse, nothing is stripped (reference autoscaler.py:52)
        self.redis_keys = dict.fromkeys(queues.split(queue_delim), 0)
        self.redis_client = redis_client
        self.logger = logging.getLogger(str(self.__class__.__name__))
        self.completed_statuses = {'done', 'failed'}
        self.managed_resource_types = {'deployment', 'job'}

    # ------------------------------------------------------------------ tally
    def _scan_term(self, names):
        """SCAN-count term of every queue: from the client's snapshot when it hands one out
        (offline replay), else from ONE streamed scan of the live server."""
        # looked up on the CLASS: RedisClient.__getattr__ (reference redis.py:163-176) answers
        # every attribute name with a command wrapper, so an instance lookup proves nothing
        taker = getattr(type(self.redis_client), 'key_snapshot', None)
        if callable(taker):
            return _snapshot.tally_snapshot(self.redis_client.key_snapshot(), names)
        return _snapshot.tally_live(self.redis_client, names, match='processing-*', count=1000)

    def tally_queues(self):
        """Update counts of all redis queues (reference autoscaler.py:60-77):
        ``LLEN q`` + number of ``processing-<q>:*`` keys, the latter counted for
        all queues by one pass of the match kernel over a key snapshot."""
        start = timeit.default_timer()
        names = list(self.redis_keys)
        waiting = []
        for q in names:
            self.logger.debug('Tallying items in queue `%s`.', q)
            waiting.append(self.redis_client.llen(q))
        in_progress = self._scan_term(names)
        for q, num_items, num_in_progress in zip(names, waiting, in_progress):
            self.redis_keys[q] = num_items + num_in_progress
        self.logger.debug('Finished tallying redis keys in %s seconds.',
                          timeit.default_timer() - start)
        self.logger.info('In-progress or new redis keys: %s', self.redis_keys)

    # ------------------------------------------------------------- kubernetes
    def get_apps_v1_client(self):
        """Returns Kubernetes API Client for AppsV1Api"""
        k = _kubernetes()
        k.config.load_incluster_config()
        return k.client.AppsV1Api()

    def get_batch_v1_client(self):
        """Returns Kubernetes API Client for BatchV1Api"""
        k = _kubernetes()
        k.config.load_incluster_config()
        return k.client.BatchV1Api()

    def _kube_call(self, client_getter, method, *args):
        """Call ``method`` on a fresh client; log and re-raise ApiException
        (reference autoscaler.py:89-151)."""
        try:
            return getattr(client_getter(), method)(*args)
        except Exception as err:
            if _is_api_exception(err):
                self.logger.error('%s when calling `%s`: %s', type(err).__name__, method, err)
            raise

    def list_namespaced_deployment(self, namespace):
        t = timeit.default_timer()
        response = self._kube_call(self.get_apps_v1_client, 'list_namespaced_deployment', namespace)
        self.logger.debug('Found %s deployments in namespace `%s` in %s seconds.',
                          len(response.items), namespace, timeit.default_timer() - t)
        return response.items

    def list_namespaced_job(self, namespace):
        t = timeit.default_timer()
        response = self._kube_call(self.get_batch_v1_client, 'list_namespaced_job', namespace)
        self.logger.debug('Found %s jobs in namespace `%s` in %s seconds.',
                          len(response.items), namespace, timeit.default_timer() - t)
        return response.items

    def patch_namespaced_deployment(self, name, namespace, body):
        t = timeit.default_timer()
        response = self._kube_call(self.get_apps_v1_client, 'patch_namespaced_deployment',
                                   name, namespace, body)
        self.logger.debug('Patched deployment `%s` in namespace `%s` with body `%s` in %s '
                          'seconds.', name, namespace, body, timeit.default_timer() - t)
        return response

    def patch_namespaced_job(self, name, namespace, body):
        t = timeit.default_timer()
        response = self._kube_call(self.get_batch_v1_client, 'patch_namespaced_job',
                                   name, namespace, body)
        self.logger.debug('Patched job `%s` in namespace `%s` with body `%s` in %s seconds.',
                          name, namespace, body, timeit.default_timer() - t)
        return response

    def get_current_pods(self, namespace, resource_type, name, only_running=False):
        """Number of pods of the resource: ``spec.replicas`` (or
        ``status.available_replicas``) of a deployment, ``spec.parallelism`` of
        a job; 0 if absent or None (reference autoscaler.py:153-195)."""
        if resource_type not in self.managed_resource_types:
            raise ValueError('`resource_type` must be one of {}. Got {}.'.format(
                self.managed_resource_types, resource_type))
        found = 0
        if resource_type == 'deployment':
            for d in self.list_namespaced_deployment(namespace):
                if d.metadata.name == name:
                    found = d.status.available_replicas if only_running else d.spec.replicas
                    self.logger.debug('Deployment %s has %s pods', name, found)
                    break
        else:
            for j in self.list_namespaced_job(namespace):
                if j.metadata.name == name:
                    found = j.spec.parallelism
                    break
        return int(found) if found is not None else 0

    # --------------------------------------------------------------- decision
    def clip_pod_count(self, desired_pods, min_pods, max_pods, current_pods):
        """Clamp to [min, max] (upper bound tested first) and never scale down
        to a non-zero count (reference autoscaler.py:197-213)."""
        clipped = engine.decide([desired_pods], None, [min_pods], [max_pods], [current_pods],
                                mode=1)[0]
        if clipped != desired_pods:
            self.logger.debug('Clipped pods from %s to %s', desired_pods, clipped)
        return clipped

    def get_desired_pods(self, key, keys_per_pod, min_pods, max_pods, current_pods):
        """``redis_keys[key] // keys_per_pod`` clipped (reference autoscaler.py:215-219)."""
        return engine.decide([self.redis_keys[key]], [keys_per_pod], [min_pods], [max_pods],
                             [current_pods], mode=0)[0]

    def _get_desired_pods_is_stock(self):
        return ('get_desired_pods' not in self.__dict__ and
                type(self).get_desired_pods is Autoscaler.get_desired_pods)

    def scale_resource(self, desired_pods, current_pods, resource_type, namespace, name):
        if resource_type not in self.managed_resource_types:
            raise ValueError('Cannot scale resource type: %s' % resource_type)
        if desired_pods == current_pods:
            return None  # no scaling action is required
        if resource_type == 'job':
            self.patch_namespaced_job(name, namespace, {'spec': {'parallelism': desired_pods}})
        else:
            self.patch_namespaced_deployment(name, namespace, {'spec': {'replicas': desired_pods}})
        self.logger.info('Successfully scaled %s `%s` in namespace `%s` from %s to %s pods.',
                         resource_type, name, namespace, current_pods, desired_pods)
        return True

    def scale(self, namespace, resource_type, name, min_pods=0, max_pods=1, keys_per_pod=1):
        """One control-loop step (reference autoscaler.py:244-273)."""
        self.tally_queues()
        self.logger.debug('Scaling %s `%s.%s`.', resource_type, namespace, name)
        current_pods = self.get_current_pods(namespace, resource_type, name)

        names = list(self.redis_keys)
        n = len(names)
        if self._get_desired_pods_is_stock():
            # every queue's decision in one launch, then the clip of their sum
            per_queue = engine.decide([self.redis_keys[q] for q in names], [keys_per_pod] * n,
                                      [min_pods] * n, [max_pods] * n, [current_pods] * n, mode=0)
        else:
            # a subclass or a test replaced get_desired_pods: call it per key, as the
            # reference does (autoscaler.py:254-257)
            per_queue = [self.get_desired_pods(q, keys_per_pod, min_pods, max_pods, current_pods)
                         for q in names]
        desired_pods = self.clip_pod_count(sum(per_queue), min_pods, max_pods, current_pods)

        self.logger.debug('%s `%s` in namespace `%s` has a current state of %s pods and a '
                          'desired state of %s pods.', str(resource_type).capitalize(), name,
                          namespace, current_pods, desired_pods)
        try:
            self.scale_resource(desired_pods, current_pods, resource_type, namespace, name)
        except Exception as err:
            if not _is_api_exception(err):
                raise
            self.logger.warning('Failed to scale %s `%s.%s` due to %s: %s', resource_type,
                                namespace, name, type(err).__name__, err)

    # ----------------------------------------------------------------- replay
    def replay(self, llen_trace, policies, inprog=None, current_pods=0,
               resource_type='deployment', want_dense=False, patch_fail=None, startup_delay=0):
        """Run ``scale`` over a whole trace for this scaler's queues (ONE
        resource) and every policy, on the GPU.  See autoscaler.replay.replay_one.
        ``startup_delay`` > 0 replays ``scale`` as if it read the pods with
        ``only_running=True`` and a replica took that many ticks to become available."""
        from autoscaler import replay as _replay
        return _replay.replay_one(list(self.redis_keys), llen_trace, policies, inprog=inprog,
                                  current_pods=current_pods, resource_type=resource_type,
                                  managed=self.managed_resource_types, want_dense=want_dense,
                                  patch_fail=patch_fail, startup_delay=startup_delay)
