"""Download cache for pretrained archives (reference: file_utils.py).

Same observable contract as the reference's HuggingFace-0.x cache -- a URL is stored under
``$PYTORCH_PRETRAINED_BERT_CACHE`` (default ``~/.pytorch_pretrained_bert``) as
``sha256(url)[.sha256(etag)]`` next to a ``.json`` file holding ``{"url", "etag"}``
(file_utils.py:34-49, 200-226) -- so a cache directory filled by the reference is reused as is.
Differences, all deliberate:

* nothing heavy is imported at module import time (the reference hard-imports ``boto3`` and
  ``requests``, which made ``bert_modeling`` unimportable without them): http(s) goes
  through ``urllib`` from the standard library, ``s3://`` imports ``boto3`` only when such a URL is
  actually requested;
* downloads are written to a temporary file in the cache directory and renamed into place, so an
  interrupted download never leaves a truncated entry;
* ``resolve_archive`` is the piece ``BertPreTrainedModel.from_pretrained`` needs: local directory,
  local ``.tar.gz`` or URL -> a directory containing ``bert_config.json`` + ``pytorch_model.bin``.

There is no network in the build/test environment: http(s)/s3 are exercised with ``file://`` URLs
in the tests, which take the same code path after the opener.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import tarfile
import tempfile
from typing import Optional, Set, Tuple
from urllib.parse import urlparse

CONFIG_NAME = 'bert_config.json'
WEIGHTS_NAME = 'pytorch_model.bin'
TF_WEIGHTS_NAME = 'model.ckpt'


def default_cache_dir() -> str:
    return os.environ.get('PYTORCH_PRETRAINED_BERT_CACHE',
                          os.path.join(os.path.expanduser('~'), '.pytorch_pretrained_bert'))


def url_to_filename(url: str, etag: Optional[str] = None) -> str:
    """Repeatable cache file name: hex sha256 of the URL, plus '.' + hex sha256 of the etag when given."""
    name = hashlib.sha256(url.encode('utf-8')).hexdigest()
    if etag:
        name += '.' + hashlib.sha256(etag.encode('utf-8')).hexdigest()
    return name


def filename_to_url(filename: str, cache_dir: Optional[str] = None) -> Tuple[str, Optional[str]]:
    """(url, etag) recorded for a cache entry; EnvironmentError if the entry or its metadata is missing."""
    cache_dir = str(cache_dir or default_cache_dir())
    path = os.path.join(cache_dir, filename)
    for p in (path, path + '.json'):
        if not os.path.exists(p):
            raise EnvironmentError('file {} not found'.format(p))
    with open(path + '.json', encoding='utf-8') as f:
        meta = json.load(f)
    return meta['url'], meta.get('etag')


def cached_path(url_or_filename, cache_dir: Optional[str] = None) -> str:
    """A local path for something that is either a URL (downloaded into the cache) or already a path."""
    url_or_filename = str(url_or_filename)
    scheme = urlparse(url_or_filename).scheme
    if scheme in ('http', 'https', 's3', 'file'):
        return get_from_cache(url_or_filename, cache_dir)
    if os.path.exists(url_or_filename):
        return url_or_filename
    if scheme == '':
        raise EnvironmentError('file {} not found'.format(url_or_filename))
    raise ValueError('unable to parse {} as a URL or as a local path'.format(url_or_filename))


def split_s3_path(url: str) -> Tuple[str, str]:
    """s3://bucket/key -> (bucket, key)."""
    parsed = urlparse(url)
    if not parsed.netloc or not parsed.path:
        raise ValueError('bad s3 path {}'.format(url))
    return parsed.netloc, parsed.path.lstrip('/')


def _s3_resource():
    try:
        import boto3
    except ImportError as e:   # pragma: no cover - boto3 is optional
        raise EnvironmentError('s3:// URLs need the optional boto3 package') from e
    return boto3.resource('s3')


def _probe(url: str) -> Optional[str]:
    """ETag of the remote object (None when the server gives none)."""
    scheme = urlparse(url).scheme
    if scheme == 's3':
        bucket, key = split_s3_path(url)
        try:
            return _s3_resource().Object(bucket, key).e_tag
        except Exception as e:
            raise EnvironmentError('file {} not found'.format(url)) from e
    if scheme == 'file':
        st = os.stat(urlparse(url).path)
        return '{}-{}'.format(st.st_size, int(st.st_mtime))
    import urllib.request
    req = urllib.request.Request(url, method='HEAD')
    try:
        with urllib.request.urlopen(req, timeout=30) as resp:
            if resp.status != 200:
                raise EnvironmentError('HEAD request failed for url {} with status code {}'.format(url, resp.status))
            return resp.headers.get('ETag')
    except OSError as e:
        raise EnvironmentError('HEAD request failed for url {}: {}'.format(url, e)) from e


def _fetch(url: str, out_file) -> None:
    scheme = urlparse(url).scheme
    if scheme == 's3':
        bucket, key = split_s3_path(url)
        _s3_resource().Bucket(bucket).download_fileobj(key, out_file)
        return
    import urllib.request
    with urllib.request.urlopen(url, timeout=60) as resp:
        shutil.copyfileobj(resp, out_file, length=1 << 20)


def get_from_cache(url: str, cache_dir: Optional[str] = None) -> str:
    """Path of the cached copy of ``url``, downloading it first when absent (or when its etag changed)."""
    cache_dir = str(cache_dir or default_cache_dir())
    os.makedirs(cache_dir, exist_ok=True)
    etag = _probe(url)
    path = os.path.join(cache_dir, url_to_filename(url, etag))
    if os.path.exists(path):
        return path
    fd, tmp = tempfile.mkstemp(dir=cache_dir, prefix='.download-')
    try:
        with os.fdopen(fd, 'wb') as f:
            _fetch(url, f)
        os.replace(tmp, path)
    except BaseException:
        if os.path.exists(tmp):
            os.remove(tmp)
        raise
    with open(path + '.json', 'w', encoding='utf-8') as f:
        json.dump({'url': url, 'etag': etag}, f)
    return path


def resolve_archive(name_or_path, cache_dir: Optional[str] = None, archive_map: Optional[dict] = None) -> str:
    """Directory holding ``bert_config.json`` (+ weights) for a model name, directory, archive or URL.

    ``archive_map`` maps short model names to URLs like the reference's PRETRAINED_MODEL_ARCHIVE_MAP
    (bert_modeling.py:33-41). ``.tar.gz`` archives are unpacked next to the cache entry once and reused
    (the reference unpacks into a fresh temp dir on every call and deletes it, bert_modeling.py:661-683)."""
    target = (archive_map or {}).get(str(name_or_path), str(name_or_path))
    local = cached_path(target, cache_dir=cache_dir)
    if os.path.isdir(local):
        return local
    if not tarfile.is_tarfile(local):
        raise EnvironmentError('{} is neither a directory nor a tar archive'.format(local))
    out_dir = local + '.extracted'
    if not os.path.exists(os.path.join(out_dir, CONFIG_NAME)):
        tmp = tempfile.mkdtemp(dir=os.path.dirname(local) or '.', prefix='.extract-')
        try:
            with tarfile.open(local, 'r:*') as archive:
                for member in archive.getmembers():     # refuse paths that escape the target directory
                    dest = os.path.realpath(os.path.join(tmp, member.name))
                    if not dest.startswith(os.path.realpath(tmp) + os.sep):
                        raise EnvironmentError('unsafe path in archive: {}'.format(member.name))
                try:
                    archive.extractall(tmp, filter='data')
                except TypeError:      # Python < 3.12 without the filter argument
                    archive.extractall(tmp)
            root = tmp
            entries = os.listdir(tmp)
            if CONFIG_NAME not in entries and len(entries) == 1 and os.path.isdir(os.path.join(tmp, entries[0])):
                root = os.path.join(tmp, entries[0])     # archive with a single top-level directory
            if os.path.exists(out_dir):
                shutil.rmtree(out_dir)
            shutil.move(root, out_dir)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
    return out_dir


def read_set_from_file(filename: str) -> Set[str]:
    """One item per line -> de-duplicated set (file_utils.py:229-238)."""
    with open(filename, 'r', encoding='utf-8') as f:
        return {line.rstrip() for line in f}


def get_file_extension(path: str, dot: bool = True, lower: bool = True) -> str:
    ext = os.path.splitext(path)[1]
    ext = ext if dot else ext[1:]
    return ext.lower() if lower else ext
