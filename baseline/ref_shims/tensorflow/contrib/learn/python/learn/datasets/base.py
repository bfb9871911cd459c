"""``Datasets`` tuple and the offline form of ``maybe_download`` (the files must already be in work_directory)."""
import collections
import os

Datasets = collections.namedtuple("Datasets", ["train", "validation", "test"])


def maybe_download(filename, work_directory, source_url):
    path = os.path.join(work_directory, filename)
    if not os.path.exists(path):
        raise IOError("no network in this image: %s must be staged beforehand (asked to fetch %s)" % (path, source_url))
    return path
