"""Problem/CLI spec, report writer, clock sampler."""
from .spec import Problem, parse_args, format_double, courant_line  # noqa: F401
from .report import format_report, output_name, write_report  # noqa: F401
