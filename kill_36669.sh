ssh test@127.0.0.1 "kill -9 3199"
ssh test@127.0.0.1 "kill -9 3200"
ssh test@127.0.0.1 "kill -9 3201"
ssh test@127.0.0.1 "kill -9 3202"
kill -9 2946
