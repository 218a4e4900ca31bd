"""`python -m pine_b200 scenes/x.pine` -- the `pine <file>` entry (src/cli/pine.cpp:9-52) for the PRL
subset of pine_b200/prl.py: builds the scene, renders it on the GPU, saves what the script saves."""
import sys

from . import prl


def main(argv):
    if len(argv) != 2:
        print("usage: python -m pine_b200 <scene.pine>", file=sys.stderr)
        return 2
    for sp in prl.run_file(argv[1]):
        c = sp.config
        print(f"rendered {sp.target.size[0]}x{sp.target.size[1]} spp={c.get('spp')} depth={c.get('depth')}"
              + (f" -> {c['save']}" if "save" in c else ""))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
