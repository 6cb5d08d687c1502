"""python -m contact_zones_b200 <config.json>"""
from .cli import main

if __name__ == "__main__":
    main()
