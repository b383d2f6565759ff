"""Install the UNMODIFIED reference into baseline/_ref (git-ignored, travels to the GPU box).  TEST / BENCH INFRASTRUCTURE.

    python oracle/install_reference.py

The base contract's recipe: `pip install --no-index --no-build-isolation --target baseline/_ref /root/reference`, from a
copy under /tmp because the build writes into the source tree (/root/reference is read-only) and with --no-deps because
the reference's requirements (mido, flatten_dict, python-rtmidi, ...) are neither in this image nor in the wheelhouse.
What the installed package can do WITHOUT those dependencies: `style.model` and `style.utils.pytorch` import and run on
torch alone, which is exactly the training half of the path, so `bench.py --impl reference` times the
reference's own StyleTransferModel + get_total_loss + backward + Adam on the host cores (`cpu_baseline.kind: "reference"`).
The codec half (style.data / style.midi_conversion) needs mido & co and stays on the C port (`kind: "port"`)."""
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = os.environ.get('MST_REFERENCE_ROOT', '/root/reference')
TARGET = os.path.join(ROOT, 'baseline', '_ref')


def installed():
    return os.path.exists(os.path.join(TARGET, 'style', 'model.py'))


def install(force=False):
    if installed() and not force:
        return TARGET
    if not os.path.isdir(os.path.join(REFERENCE, 'style')):
        raise RuntimeError(f'{REFERENCE} is not present: baseline/_ref can only be built in the build container')
    with tempfile.TemporaryDirectory() as tmp:
        src = os.path.join(tmp, 'reference')
        shutil.copytree(REFERENCE, src, ignore=shutil.ignore_patterns('examples', 'latex', '.git'))
        shutil.rmtree(TARGET, ignore_errors=True)
        os.makedirs(TARGET, exist_ok=True)
        wheelhouse = ['--find-links', '/opt/wheelhouse'] if os.path.isdir('/opt/wheelhouse') else []
        subprocess.check_call([sys.executable, '-m', 'pip', 'install', '--quiet', '--no-index', '--no-build-isolation', '--no-deps',
                               *wheelhouse, '--target', TARGET, src])
    assert installed()
    return TARGET


if __name__ == '__main__':
    print(install(force='--force' in sys.argv))
