"""Command line entry point -- same flags and defaults as the reference's ``main.py:13-78``."""
import argparse

from .make_aterm_images import make_aterm_image

_FLAGS = (
    ("--soltabname", "Name of soltab", str, "phase000"),
    ("--screen_type", "Type of screen", str, "tessellated"),
    ("--outroot", "Root of output images", str, ""),
    ("--bounds_deg", "Bounds list in deg", str, None),
    ("--bounds_mid_deg", "Bounds mid list in deg", str, None),
    ("--skymodel", "Filename of sky model", str, None),
    ("--solsetname", "Solset name", str, "sol000"),
    ("--padding_fraction", "Padding fraction", float, 1.4),
    ("--cellsize_deg", "Cell size in deg", float, 0.2),
    ("--smooth_deg", "Smooth scale in degree", float, 0.0),
    ("--ncpu", "Number of CPUs to use", int, 0),
)


def build_parser():
    parser = argparse.ArgumentParser(description="Make a-term images from solutions.\n",
                                     formatter_class=argparse.RawTextHelpFormatter)
    parser.add_argument("h5parmfile", help="Filename of input H5parm")
    for flag, text, kind, default in _FLAGS:
        parser.add_argument(flag, help=text, type=kind, default=default)
    return parser


def start(argv=None):
    args = build_parser().parse_args(argv)
    options = {flag.lstrip("-"): getattr(args, flag.lstrip("-")) for flag, *_ in _FLAGS}
    make_aterm_image(args.h5parmfile, **options)


if __name__ == "__main__":
    start()
