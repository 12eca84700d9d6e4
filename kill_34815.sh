ssh test@127.0.0.1 "kill -9 2362"
ssh test@127.0.0.1 "kill -9 2363"
ssh test@127.0.0.1 "kill -9 2364"
ssh test@127.0.0.1 "kill -9 2365"
kill -9 2109
