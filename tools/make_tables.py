"""Derive the compact index tables the hot path needs from a pokeai checkout.

The reference locates its tables through ``pokeai.util.DATASET_DIR``
(/root/reference/pokeai/util.py:9-10) and reads four JSON files:
``all_pokemons.json`` / ``all_moves_with_hiddenpower_type.json`` /
``all_items.json`` (choice_to_vec.py:18-26: list order == one-hot row order)
and ``pokedex.json`` (dex.py:10-11: display name -> entry with ``types``).

Only the *index order* and the species -> type list matter to the encoder, so
this script writes one small JSON (``pokeai_b200/data/tables.json``) holding
exactly that.  Run it in the build container where the reference is mounted:

    python tools/make_tables.py /root/reference/data/dataset
"""
import json
import sys
from pathlib import Path

# state_feature_extractor.py:9-27 -- order of the 17-way type multi-hot
POKE_TYPES = ["Normal", "Fighting", "Flying", "Poison", "Ground", "Rock", "Bug", "Ghost", "Steel",
              "Fire", "Water", "Grass", "Electric", "Psychic", "Ice", "Dragon", "Dark"]


def main(dataset_dir: str) -> None:
    d = Path(dataset_dir)
    load = lambda n: json.loads(d.joinpath(n).read_text(encoding="utf-8"))
    species_ids = load("all_pokemons.json")
    moves = load("all_moves_with_hiddenpower_type.json")
    items = load("all_items.json")
    pokedex = load("pokedex.json")
    t2n = {t: i for i, t in enumerate(POKE_TYPES)}
    species = []
    for sid in species_ids:
        e = pokedex[sid]
        assert e["id"] == sid
        species.append({"id": sid, "name": e["name"], "types": [t2n[t] for t in e["types"]]})
    # every pokedex entry must be addressable by index (u8 in the packed record)
    assert set(pokedex) == set(species_ids) and len(species_ids) < 256
    out = {
        "source": "select766/pokeai data/dataset (gen2customgame tables)",
        "poke_types": POKE_TYPES,
        "species": species,
        "moves": moves,
        "items": items,
    }
    dst = Path(__file__).resolve().parents[1] / "pokeai_b200" / "data" / "tables.json"
    dst.write_text(json.dumps(out, ensure_ascii=False, separators=(",", ":")), encoding="utf-8")
    print(f"wrote {dst}: {len(species)} species, {len(moves)} moves, {len(items)} items")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "/root/reference/data/dataset")
