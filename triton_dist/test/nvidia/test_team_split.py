"""Teams (mirrors reference test/nvidia/test_team_split.py: EP x PP split of the world, my_pe / n_pes / translate_pe, a
barrier and a put inside a sub-team).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_team_split --ep-size 4"""
import torch

from triton_dist.runtime import get_heap, team_world
from triton_dist.test.common import all_ok, base_parser, finish, setup
from triton_dist.utils import dist_print, nvshmem_barrier_all_on_stream, symm_tensor

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--ep-size", type=int, default=0, help="default: world / 2 (or 1)")
    args = p.parse_args()
    group, rank, world = setup()
    ep = args.ep_size or max(world // 2, 1)
    assert world % ep == 0
    pp = world // ep
    tw = team_world(world, rank)
    ep_team, pp_team = tw.split_2d(ep)
    ok = ep_team.n_pes() == ep and pp_team.n_pes() == pp
    ok = ok and ep_team.my_pe() == rank % ep and pp_team.my_pe() == rank // ep
    ok = ok and ep_team.ranks() == [rank // ep * ep + i for i in range(ep)]
    ok = ok and pp_team.ranks() == [rank % ep + i * ep for i in range(pp)]
    # translate: my EP rank -> world -> back; a rank of another row is not in my EP team
    ok = ok and ep_team.translate_pe(ep_team.my_pe(), tw) == rank and tw.translate_pe(rank, ep_team) == ep_team.my_pe()
    if pp > 1:
        other = (rank + ep) % world
        ok = ok and tw.translate_pe(other, ep_team) == -1 and tw.translate_pe(other, pp_team) == other // ep
    # strided split: even world ranks
    if world >= 2:
        ev = tw.split_strided(0, 2, world // 2)
        ok = ok and ev.ranks() == list(range(0, world, 2)) and ev.my_pe() == (rank // 2 if rank % 2 == 0 else -1)
    # data movement inside the EP team: every member writes its team rank into every member's slot, team barrier
    heap = get_heap()
    box = symm_tensor((ep,), torch.int32)
    flags = symm_tensor((ep,), torch.int32)
    nvshmem_barrier_all_on_stream()
    for it in range(1, 4):
        me = ep_team.my_pe()
        if torch.cuda.is_available():
            for r in range(ep):
                heap.peer_view(box, ep_team.world_rank(r))[me:me + 1].fill_(100 * it + me)
            ep_team.barrier(flags, it)
            torch.cuda.synchronize()
            ok = ok and box.tolist() == [100 * it + r for r in range(ep)]
        else:
            ep_team.barrier(flags, it)
        nvshmem_barrier_all_on_stream()
    dist_print(f"rank {rank}: EP team {ep_team.my_pe()}/{ep_team.n_pes()} {ep_team.ranks()}, PP team {pp_team.my_pe()}/{pp_team.n_pes()} "
               f"{pp_team.ranks()}", allowed_ranks="all")
    finish(all_ok(ok))
