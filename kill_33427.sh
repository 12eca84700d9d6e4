ssh test@127.0.0.1 "kill -9 5726"
ssh test@127.0.0.1 "kill -9 5727"
ssh test@127.0.0.1 "kill -9 5728"
kill -9 5044
